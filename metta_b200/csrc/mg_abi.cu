// mg_abi.cu -- the C-ABI of libmettagrid_b200.so (declared in include/mettagrid_b200.h).
//
// Plain pointers and sizes only. A batch of environments is an opaque handle; observation,
// reward, terminal, truncation and action buffers are caller-owned device memory bound with
// mg_set_buffers, exactly like MettaGrid::set_buffers binds caller-owned numpy arrays
// (bindings/mettagrid_c.cpp:721-738). Errors come back as non-zero status codes with a message
// in mg_last_error(); nothing here falls back to the CPU.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "mg_kernels.cuh"

static thread_local std::string g_err;

#define MG_MAX_CHUNKS 16
// Dynamic shared memory of the step kernel can hold a per-CTA copy of the lowered config (up to this
// many bytes). Measured on B200 (arena, 4096 envs): the copy costs more than the L1 hits it buys
// (step kernel 128.6 us with it, 120.7 us without), so it is off; -DMG_CFG_SMEM_LIMIT=98304 turns it on.
#ifndef MG_CFG_SMEM_LIMIT
#define MG_CFG_SMEM_LIMIT 0
#endif
#ifndef MG_DEFAULT_STEP_MODE
#define MG_DEFAULT_STEP_MODE 0
#endif
// observation CTAs per SM that run beside the step kernel in queue mode (the rest of the SM stays with
// the step kernel, so it always makes progress), and that join once the step kernel has finished
#ifndef MG_QUEUE_OBS_CTAS_EARLY
#define MG_QUEUE_OBS_CTAS_EARLY 4
#endif
#ifndef MG_QUEUE_OBS_CTAS_LATE
#define MG_QUEUE_OBS_CTAS_LATE 4
#endif
#define MG_OK 0
#define MG_ERR_ARG 1
#define MG_ERR_CUDA 2
#define MG_ERR_STATE 3

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t err__ = (expr);                                                                 \
    if (err__ != cudaSuccess)                                                                   \
      return fail(MG_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(err__));          \
  } while (0)

struct MgEnvs {
  MgBatch b;
  int device;
  std::vector<int32_t> cfg_host;
  std::vector<void*> allocs;
  long long launches;
  bool maps_loaded, buffers_bound, owns_buffers;
  // host-buffer path (mg_step_host): device mirrors + pinned staging
  uint8_t* h_pinned;
  size_t h_pinned_bytes;
  cudaStream_t stream;
  // the step is launched as MG_CHUNKS independent slices of envs on internal streams, so that the
  // observation kernel of one slice overlaps the (latency-bound) step kernel of the others; the whole
  // fork/join pattern is captured once in a CUDA graph and replayed
  cudaStream_t chunk_stream[MG_MAX_CHUNKS], capture_stream;
  cudaEvent_t ev_fork, ev_join[MG_MAX_CHUNKS];
  int chunks;
  int chunks_host;  // slices of the step graph that carries the host export (more of them: the exports of early slices overlap the later slices)
  // 1: step and observation kernels run side by side, finished envs handed over through a queue
  // (enqueue_queue_step); 0: slices of envs, each slice's observation kernel behind its step kernel
  int step_mode;
  int sm_count;
  int q_early, q_late;  // observation CTAs per SM beside / behind the step kernel in queue mode
  cudaStream_t prio_hi, prio_lo;  // flag mode: the step kernel's stream (its pending CTAs go first) and the observation kernel's
  // mg_step_host_begin .. mg_step_host_end: what _end still has to do
  struct HostPending {
    bool active = false, staged = false;
    uint8_t* obs = nullptr; float* rewards = nullptr; uint8_t* terminals = nullptr; uint8_t* truncations = nullptr;
    size_t NA = 0, obs_bytes = 0, off_obs = 0, off_rew = 0, off_term = 0, off_trunc = 0;
  } host_pending;
  bool graph_dirty;
  long long graph_launches, hgraph_launches;  // kernels per replay of the two graphs
  cudaGraph_t graph;
  cudaGraphExec_t graph_exec;
  uint8_t* d_mask;
  uint32_t* d_seeds;   // mg_reset: seeds and map indices of the envs being reset (patched into EnvRec on the device)
  int32_t* d_map_idx;
  int32_t* d_tmp;  // small scratch for set_inventory
  // library-owned caller buffers (mg_set_buffers with all NULLs), kept so that a later all-NULL call rebinds them
  uint8_t *own_obs, *own_term, *own_trunc;
  float* own_rew;
  int32_t* own_act;
  // mg_step_host with page-locked buffers, export mode 1: only the part of each row that changed
  int host_export_mode;
  const void* host_obs_last;      // the host buffer whose contents d_prev_counts describes (nullptr = none)
  uint16_t* d_prev_counts;        // [N*A] tokens the host row holds
  uint8_t* d_shadow;              // [N*A][T*3] what the host buffer holds (allocated when export mode 1 is selected)
  unsigned long long* d_export_bytes;
  unsigned long long* h_export_bytes;  // pinned mirror of the counter
  unsigned long long h2d_total, d2h_total;
  // the step graph with the export kernels of every slice in it, for one set of host buffers
  cudaGraph_t hgraph;
  cudaGraphExec_t hgraph_exec;
  bool hgraph_dirty;
  MgHostTargets hgraph_targets;
};

static void mark_graphs_dirty(MgEnvs* e) { e->graph_dirty = true; e->hgraph_dirty = true; }

template <typename T>
static int dev_alloc(MgEnvs* e, T** out, size_t count, bool zero = true) {
  void* p = nullptr;
  size_t bytes = count * sizeof(T);
  if (bytes == 0) bytes = 16;
  CUDA_TRY(cudaMalloc(&p, bytes));
  if (zero) CUDA_TRY(cudaMemset(p, 0, bytes));
  e->allocs.push_back(p);
  *out = reinterpret_cast<T*>(p);
  return MG_OK;
}

// Queue mode only: did an observation CTA give up waiting for the step kernel? (checked after a sync)
static int check_queue_error(MgEnvs* e, const char* who) {
  if (e->step_mode == 0 || e->b.qctl == nullptr) return MG_OK;
  MgQueueCtl ctl;
  CUDA_TRY(cudaMemcpy(&ctl, e->b.qctl, sizeof(ctl), cudaMemcpyDeviceToHost));
  if (ctl.error != 0)
    return fail(MG_ERR_CUDA, std::string(who) + ": an observation CTA gave up waiting for the step kernel (queue mode); results are incomplete");
  return MG_OK;
}

// systems/packed_coordinate.hpp:74-143: offsets by Manhattan distance, dr ascending, -dc before +dc
static std::vector<uint8_t> make_pattern(int oh, int ow) {
  std::vector<uint8_t> out;
  const int hr = oh / 2, wr = ow / 2;
  for (int d = 0; d <= hr + wr; d++)
    for (int dr = -d; dr <= d; dr++) {
      const int adc = d - (dr < 0 ? -dr : dr);
      for (int s = 0; s < (adc == 0 ? 1 : 2); s++) {
        const int dc = s == 0 ? -adc : adc;
        if (dr < -hr || dr > hr || dc < -wr || dc > wr) continue;
        out.push_back((uint8_t)(((dr + hr) << 4) | ((dc + wr) & 0xF)));
      }
    }
  return out;
}

extern "C" {

const char* mg_last_error(void) { return g_err.c_str(); }
int mg_abi_version(void) { return MG_CONFIG_VERSION; }

int mg_create(const int32_t* cfg, int64_t n_words, int32_t num_envs, int32_t device, void** out) {
  if (!cfg || !out || n_words < MGC_HEADER_WORDS) return fail(MG_ERR_ARG, "mg_create: null or short config");
  if (cfg[MGC_MAGIC] != MG_CONFIG_MAGIC || cfg[MGC_VERSION] != MG_CONFIG_VERSION)
    return fail(MG_ERR_ARG, "mg_create: config blob has the wrong magic/version");
  if (cfg[MGC_TOTAL_WORDS] != n_words) return fail(MG_ERR_ARG, "mg_create: config length mismatch");
  if (num_envs <= 0) return fail(MG_ERR_ARG, "mg_create: num_envs must be positive");
  if (cfg[MGC_NUM_AGENTS] > MG_MAX_AGENTS)
    return fail(MG_ERR_ARG, "mg_create: more than " + std::to_string(MG_MAX_AGENTS) + " agents per env");
  if (cfg[MGC_OBS_H] * cfg[MGC_OBS_W] > 256) return fail(MG_ERR_ARG, "mg_create: observation window too large");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(MG_ERR_CUDA, "mg_create: no CUDA device (this library has no CPU path)");
  if (device < 0 || device >= ndev) return fail(MG_ERR_ARG, "mg_create: bad device ordinal");
  CUDA_TRY(cudaSetDevice(device));
  // inv_update <-> inv_enforce_all_limits recurse (bounded depth 8); give device threads room for it
  CUDA_TRY(cudaDeviceSetLimit(cudaLimitStackSize, 8192));
  MgEnvs* e = new MgEnvs();
  memset(&e->b, 0, sizeof(MgBatch));
  e->device = device;
  e->launches = 0;
  e->maps_loaded = e->buffers_bound = e->owns_buffers = false;
  e->h_pinned = nullptr;
  e->h_pinned_bytes = 0;
  e->stream = nullptr;
  e->chunks = 1;
  mark_graphs_dirty(e);
  e->graph = nullptr;
  e->graph_exec = nullptr;
  e->ev_fork = nullptr;
  e->capture_stream = nullptr;
  for (int i = 0; i < MG_MAX_CHUNKS; i++) { e->chunk_stream[i] = nullptr; e->ev_join[i] = nullptr; }
  e->prio_hi = e->prio_lo = nullptr;
  e->cfg_host.assign(cfg, cfg + n_words);
  MgBatch& b = e->b;
  b.num_envs = num_envs;
  b.A = cfg[MGC_NUM_AGENTS];
  b.H = cfg[MGC_MAP_H];
  b.W = cfg[MGC_MAP_W];
  b.T = cfg[MGC_NUM_TOKENS];
  b.R = cfg[MGC_NUM_RESOURCES];
  b.S = cfg[MGC_NUM_AGENT_STATS];
  b.GS = cfg[MGC_NUM_GAME_STATS];
  b.max_asm = cfg[MGC_MAX_ASSEMBLERS] > 0 ? cfg[MGC_MAX_ASSEMBLERS] : 1;
  b.max_chest = cfg[MGC_MAX_CHESTS] > 0 ? cfg[MGC_MAX_CHESTS] : 1;
  b.has_clipper = (cfg + cfg[MGC_OFF_CLIPPER])[MG_CLIP_ENABLED];
  b.obs_cells = cfg[MGC_OBS_H] * cfg[MGC_OBS_W];
  b.obs_hr = cfg[MGC_OBS_H] >> 1;
  b.obs_wr = cfg[MGC_OBS_W] >> 1;
  b.feat_cooldown = cfg[MGC_FEAT_COOLDOWN];
  b.grid_stride = ((b.H * b.W + 63) / 64) * 64;  // whole 128-byte lines per env (see mg_fused_kernel)
  b.auto_reset = 1;
  b.map_stride = 0;
  b.env_begin = 0;
  b.env_end = num_envs;
  const size_t N = (size_t)num_envs, A = (size_t)b.A;
  int32_t* d_cfg = nullptr;
  uint8_t* d_pat = nullptr;
  int rc;
#define TRY(x) if ((rc = (x)) != MG_OK) { return rc; }
  // behind the blob: for every INVCFG record and limit, which items share the limit and which items
  // modify it, one bit per item, so that the inventory code walks set bits instead of all resources
  std::vector<int32_t> cfg_ext(cfg, cfg + n_words);
  for (int ic = 0; ic < cfg[MGC_NUM_INVCFG]; ic++) {
    const int32_t* rec = cfg + cfg[MGC_OFF_INVCFG] + (size_t)ic * MG_INVCFG_WORDS;
    int32_t members[MG_MAX_LIMITS] = {0}, mods[MG_MAX_LIMITS] = {0};
    for (int l = 0; l < MG_MAX_LIMITS; l++)
      for (int i = 0; i < MG_MAX_RESOURCES; i++) {
        if (rec[MG_INVCFG_ITEM_LIMIT + i] == l) members[l] |= 1 << i;
        if (rec[MG_INVCFG_LIMIT_MOD + l * MG_RES_STRIDE + i] != 0) mods[l] |= 1 << i;
      }
    cfg_ext.insert(cfg_ext.end(), members, members + MG_MAX_LIMITS);
    cfg_ext.insert(cfg_ext.end(), mods, mods + MG_MAX_LIMITS);
  }
  TRY(dev_alloc(e, &d_cfg, cfg_ext.size() + 4));
  CUDA_TRY(cudaMemcpy(d_cfg, cfg_ext.data(), sizeof(int32_t) * cfg_ext.size(), cudaMemcpyHostToDevice));
  b.cfg = d_cfg;
  b.cfg_words = (int32_t)cfg_ext.size();
  b.inv_masks_off = (int32_t)n_words;
  {
    const int32_t* res = cfg + cfg[MGC_OFF_ACTION_RES];
    for (int k = 0; k < 3; k++) {
      b.req_mask[k] = b.con_mask[k] = 0u;
      for (int i = 0; i < MG_MAX_RESOURCES; i++) {
        if ((res[k * MG_ACTRES_WORDS + MG_ACTRES_REQUIRED + i] & 0xFFFF) != 0) b.req_mask[k] |= 1u << i;
        if ((res[k * MG_ACTRES_WORDS + MG_ACTRES_CONSUMED + i] & 0xFFFF) != 0) b.con_mask[k] |= 1u << i;
      }
    }
    b.any_damage = b.reward_stat_prefetch = b.reward_game_stats = 0;
    for (int g = 0; g < cfg[MGC_NUM_GROUPS]; g++) {
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + (size_t)g * MG_GRP_WORDS;
      if (grp[MG_GRP_DAMAGE_ENABLED] != 0) b.any_damage = 1;
      for (int i = 0; i < grp[MG_GRP_NUM_REWARDS]; i++) {
        const int32_t* rw = grp + MG_GRP_REWARDS + i * MG_RWD_WORDS;
        const int slot = rw[MG_RWD_AGENT_STAT];
        if (slot >= MG_HOT_STATS && !(slot >= MG_AS_AMOUNT && slot < MG_AS_AMOUNT + MG_MAX_RESOURCES)) b.reward_stat_prefetch = 1;
        if (rw[MG_RWD_GAME_STAT] >= 0) b.reward_game_stats = 1;
      }
    }
  }
  b.any_chest_search = 0;
  for (int c = 1; c < cfg[MGC_NUM_CELLTYPES] && c < MG_MAX_CELLTYPES; c++) {
    const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c * MG_CELL_WORDS;
    if (cc[MG_CELL_KIND] == MG_KIND_ASSEMBLER && cfg[cfg[MGC_OFF_ASM_TYPES] + cc[MG_CELL_PARAM] * MG_ASM_WORDS + MG_ASM_CHEST_DISTANCE] > 0) b.any_chest_search = 1;
  }
  {  // one line per map cell code with everything a use of a one-protocol assembler reads (assembler_use_warp)
    std::vector<int32_t> fast((size_t)MG_MAX_CELLTYPES * MG_AF_WORDS, 0);
    for (int c = 1; c < cfg[MGC_NUM_CELLTYPES] && c < MG_MAX_CELLTYPES; c++) {
      const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c * MG_CELL_WORDS;
      if (cc[MG_CELL_KIND] != MG_KIND_ASSEMBLER) continue;
      const int32_t* ty = cfg + cfg[MGC_OFF_ASM_TYPES] + cc[MG_CELL_PARAM] * MG_ASM_WORDS;
      int32_t* f = &fast[(size_t)c * MG_AF_WORDS];
      f[MG_AF_MAX_USES] = ty[MG_ASM_MAX_USES];
      f[MG_AF_PARTIAL] = ty[MG_ASM_ALLOW_PARTIAL];
      f[MG_AF_CHEST_DISTANCE] = ty[MG_ASM_CHEST_DISTANCE];
      if (ty[MG_ASM_NUM_PROTOCOLS] != 1) continue;
      const int pi = ty[MG_ASM_FIRST_PROTOCOL];
      const int32_t* p = cfg + cfg[MGC_OFF_PROTOCOLS] + pi * MG_PROTO_WORDS;
      // the one protocol must be what get_current_protocol finds for a lone actor of any vibe: key 0, min_agents <= 1
      if (p[MG_PROTO_KEY_LO] != 0 || p[MG_PROTO_KEY_HI] != 0 || p[MG_PROTO_MIN_AGENTS] > 1) continue;
      const int32_t* lin = p + MG_PROTO_INPUTS;
      const int32_t* lout = p + MG_PROTO_OUTPUTS;
      if (lin[MG_LIST_N] > 4 || lout[MG_LIST_N] > 4) continue;
      f[MG_AF_COOLDOWN] = p[MG_PROTO_COOLDOWN] & 0xFFFF;
      f[MG_AF_N_IN] = lin[MG_LIST_N];
      f[MG_AF_N_OUT] = lout[MG_LIST_N];
      f[MG_AF_PROTOCOL] = pi;
      for (int i = 0; i < lin[MG_LIST_N]; i++) { f[MG_AF_IN_ITEM + i] = lin[MG_LIST_ITEM(i)]; f[MG_AF_IN_AMOUNT + i] = lin[MG_LIST_VAL(i)] & 0xFFFF; }
      for (int i = 0; i < lout[MG_LIST_N]; i++) { f[MG_AF_OUT_ITEM + i] = lout[MG_LIST_ITEM(i)]; f[MG_AF_OUT_AMOUNT + i] = lout[MG_LIST_VAL(i)] & 0xFFFF; }
      f[MG_AF_VALID] = 1;
    }
    int32_t* d_fast;
    TRY(dev_alloc(e, &d_fast, fast.size(), false));
    CUDA_TRY(cudaMemcpy(d_fast, fast.data(), fast.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    b.asm_fast = d_fast;
  }
  std::vector<uint8_t> pat = make_pattern(cfg[MGC_OBS_H], cfg[MGC_OBS_W]);
  pat.resize(((pat.size() + 255) / 256 + 1) * 256, 0xFF);  // padded: lanes read whole runs of cells
  TRY(dev_alloc(e, &d_pat, pat.size()));
  CUDA_TRY(cudaMemcpy(d_pat, pat.data(), pat.size(), cudaMemcpyHostToDevice));
  b.pattern = d_pat;
  {
    std::vector<int32_t> off(pat.size(), -1);
    for (size_t i = 0; i < pat.size(); i++)
      if (pat[i] != 0xFF) off[i] = (int32_t)(pat[i] >> 4) * b.W + (int32_t)(pat[i] & 0xF);
    int32_t* d_off;
    TRY(dev_alloc(e, &d_off, off.size(), false));
    CUDA_TRY(cudaMemcpy(d_off, off.data(), off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    b.pattern_off = d_off;
    const int cpl = (b.obs_cells + 31) / 32;
    const int cpl_t = cpl <= 1 ? 1 : cpl <= 2 ? 2 : cpl <= 4 ? 4 : 8;
    b.obs_slot = (b.T * 3 + 16 + 15) & ~15;
    b.obs_warp_bytes = b.obs_slot + MG_OBS_DESC_BYTES(cpl_t);
  }
  TRY(dev_alloc(e, &b.env, N));
  TRY(dev_alloc(e, &b.grid, N * (size_t)b.grid_stride));
  TRY(dev_alloc(e, &b.agents, N * A));
  TRY(dev_alloc(e, &b.asms, N * (size_t)b.max_asm));
  TRY(dev_alloc(e, &b.chests, N * (size_t)b.max_chest));
  TRY(dev_alloc(e, &b.agent_stats, N * A * (size_t)b.S));
  TRY(dev_alloc(e, &b.hot_stats, N * A * (size_t)MG_HOT_STATS));
  TRY(dev_alloc(e, &b.game_stats, N * (size_t)b.GS));
  TRY(dev_alloc(e, &b.rng, N * MT_N));
  TRY(dev_alloc(e, &b.rng_seeded, N * MT_N, false));
  if (b.has_clipper) TRY(dev_alloc(e, &b.clip_rng, N * MT_N));
  {
    // static token templates (see mg_obs.cuh): cells, then protocols, then agent groups
    std::vector<uint16_t> tok;
    std::vector<uint32_t> ix;
    auto tk = [](int feature, int value) { return (uint16_t)((feature & 0xFF) | ((value & 0xFF) << 8)); };
    const int ncell = cfg[MGC_NUM_CELLTYPES], nproto = cfg[MGC_NUM_PROTOCOLS], ngroup = cfg[MGC_NUM_GROUPS];
    int max_tags = 0;
    for (int c = 0; c < ncell; c++) {
      const int32_t* cc = cfg + cfg[MGC_OFF_CELLTYPES] + c * MG_CELL_WORDS;
      const size_t off = tok.size();
      if (c > 0 && cc[MG_CELL_KIND] != MG_KIND_AGENT) {
        for (int i = 0; i < cc[MG_CELL_NUM_TAGS]; i++) tok.push_back(tk(cfg[MGC_FEAT_TAG], cc[MG_CELL_TAGS + i]));
        // walls and assemblers end with their vibe (wall.hpp:26-38, assembler.hpp:570-575); a chest's vibe comes first
        if (cc[MG_CELL_KIND] != MG_KIND_CHEST && cc[MG_CELL_VIBE] != 0) tok.push_back(tk(cfg[MGC_FEAT_VIBE], cc[MG_CELL_VIBE]));
        if (cc[MG_CELL_NUM_TAGS] > max_tags) max_tags = cc[MG_CELL_NUM_TAGS];
      }
      ix.push_back((uint32_t)(off << 8) | (uint32_t)(tok.size() - off));
    }
    b.ix_proto = (int32_t)ix.size();
    const int32_t* pf = cfg + cfg[MGC_OFF_PROTO_FEATURES];
    for (int p = 0; p < nproto; p++) {  // assembler.hpp:553-564: inputs then outputs with amount > 0
      const int32_t* pc = cfg + cfg[MGC_OFF_PROTOCOLS] + p * MG_PROTO_WORDS;
      const size_t off = tok.size();
      const int32_t* l = pc + MG_PROTO_INPUTS;
      for (int i = 0; i < l[MG_LIST_N]; i++)
        if (l[MG_LIST_VAL(i)] > 0) tok.push_back(tk(pf[l[MG_LIST_ITEM(i)]], l[MG_LIST_VAL(i)]));
      l = pc + MG_PROTO_OUTPUTS;
      for (int i = 0; i < l[MG_LIST_N]; i++)
        if (l[MG_LIST_VAL(i)] > 0) tok.push_back(tk(pf[MG_RES_STRIDE + l[MG_LIST_ITEM(i)]], l[MG_LIST_VAL(i)]));
      ix.push_back((uint32_t)(off << 8) | (uint32_t)(tok.size() - off));
    }
    b.ix_group = (int32_t)ix.size();
    for (int g = 0; g < ngroup; g++) {
      const int32_t* grp = cfg + cfg[MGC_OFF_GROUPS] + g * MG_GRP_WORDS;
      const size_t off = tok.size();
      for (int i = 0; i < grp[MG_GRP_NUM_TAGS]; i++) tok.push_back(tk(cfg[MGC_FEAT_TAG], grp[MG_GRP_TAGS + i]));
      if (grp[MG_GRP_NUM_TAGS] > max_tags) max_tags = grp[MG_GRP_NUM_TAGS];
      ix.push_back((uint32_t)(off << 8) | (uint32_t)(tok.size() - off));
    }
    tok.resize(tok.size() + 8, 0);
    uint16_t* d_tok; uint32_t* d_ix;
    TRY(dev_alloc(e, &d_tok, tok.size()));
    TRY(dev_alloc(e, &d_ix, ix.size()));
    CUDA_TRY(cudaMemcpy(d_tok, tok.data(), tok.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(d_ix, ix.data(), ix.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    b.tmpl_tok = d_tok; b.tmpl_ix = d_ix;
    // token records (format: MgBatch in mg_state.h): one power-of-two stride fits the largest list
    // any object can show: agent = group, frozen, vibe + inventory digits + tags; assembler =
    // clipped, remaining_uses + protocol inputs/outputs + tags + vibe (cooldown is added by observers)
    const int max_agent_tokens = 3 + b.R * cfg[MGC_NUM_INV_TOKENS] + max_tags;
    const int max_asm_tokens = 2 + 2 * b.R + max_tags + 1;
    const int max_tokens = max_agent_tokens > max_asm_tokens ? max_agent_tokens : max_asm_tokens;
    if (max_tokens > 254) return fail(MG_ERR_ARG, "mg_create: an object could show more than 254 tokens");
    b.rec_shift = 5;
    while ((1 << b.rec_shift) < MG_REC_TOKENS + 2 * max_tokens) b.rec_shift++;
    const int stride = 1 << b.rec_shift;
    const int base_agent = ncell, base_asm = ncell + b.A, base_chest = ncell + b.A + b.max_asm;
    b.rec_slots = base_chest + b.max_chest;
    {  // pad so that an env's records end on a 128-byte line
      const int per_line = 128 >> b.rec_shift;
      if (per_line > 1) b.rec_slots = ((b.rec_slots + per_line - 1) / per_line) * per_line;
    }
    if (b.rec_slots > 0xFFFF) return fail(MG_ERR_ARG, "mg_create: too many objects per env for the token record index");
    b.rec_kind_base_lo = (uint32_t)0 | ((uint32_t)base_agent << 16);   // grid kinds 0 static, 1 agent
    b.rec_kind_base_hi = (uint32_t)base_asm | ((uint32_t)base_chest << 16);  // 2 assembler, 3 chest
    TRY(dev_alloc(e, &b.etok, (N * (size_t)b.rec_slots) << b.rec_shift));
    TRY(dev_alloc(e, &b.gtok, N * A * (size_t)MG_GTOK_STRIDE));
    if (3 + cfg[MGC_GOBS_COMPASS] + MG_MAX_GOALS > (MG_GTOK_STRIDE - 1) / 3)
      return fail(MG_ERR_ARG, "mg_create: internal: global token record too small");
    {  // static objects as token records, indexed by cell code; copied into every env at reset
      std::vector<uint8_t> srec((size_t)ncell * stride, 0);
      for (int c = 0; c < ncell; c++) {
        const uint32_t en = ix[c];
        srec[(size_t)c * stride] = (uint8_t)(en & 0xFF);
        memcpy(&srec[(size_t)c * stride + MG_REC_TOKENS], &tok[en >> 8], (en & 0xFF) * sizeof(uint16_t));
      }
      uint8_t* d_srec;
      TRY(dev_alloc(e, &d_srec, srec.size()));
      CUDA_TRY(cudaMemcpy(d_srec, srec.data(), srec.size(), cudaMemcpyHostToDevice));
      b.static_rec = d_srec;
    }
  }
  TRY(dev_alloc(e, &b.tok_counts, N * A * 2));
  TRY(dev_alloc(e, &b.last_action, N * A));
  TRY(dev_alloc(e, &b.action_success, N * A));
  TRY(dev_alloc(e, &b.invalid_ids, N * A * MG_INVALID_WORDS));
  TRY(dev_alloc(e, &b.episode_accum, (size_t)(2 + b.S + b.GS)));
  TRY(dev_alloc(e, &e->d_mask, N));
  TRY(dev_alloc(e, &e->d_seeds, N));
  TRY(dev_alloc(e, &e->d_map_idx, N));
  TRY(dev_alloc(e, &e->d_tmp, (size_t)4 * MG_RES_STRIDE));
  CUDA_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  TRY(dev_alloc(e, &e->d_prev_counts, N * A));
#ifdef MG_TIMING
  TRY(dev_alloc(e, &b.dbg, N * 16));
#endif
  TRY(dev_alloc(e, &e->d_export_bytes, 1));
  CUDA_TRY(cudaMallocHost((void**)&e->h_export_bytes, sizeof(unsigned long long)));
  *e->h_export_bytes = 0;
#ifndef MG_SIMT_EMU
  CUDA_TRY(cudaFuncSetAttribute(mg_step_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(sizeof(StepSmem<32>) * MG_STEP_WARPS + MG_CFG_SMEM_LIMIT)));
  CUDA_TRY(cudaFuncSetAttribute(mg_step_kernel<MG_MAX_AGENTS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(sizeof(StepSmem<MG_MAX_AGENTS>) * MG_STEP_WARPS + MG_CFG_SMEM_LIMIT)));
  // Kernels whose shared-memory carve-outs differ do not share an SM: a CTA of the observation kernel would wait
  // until the SM has drained of step CTAs (and the other way round) and the two kernels of a step would only
  // overlap at SM granularity. Ask for the same carve-out for all of them (MG_SMEM_CARVEOUT: per cent of the
  // maximum; 0 = leave the choice to the driver).
  if (const char* v = getenv("MG_SMEM_CARVEOUT")) {
    const int pct = atoi(v);
    if (pct > 0) {
      CUDA_TRY(cudaFuncSetAttribute(mg_step_kernel<32>, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
      CUDA_TRY(cudaFuncSetAttribute(mg_step_kernel<MG_MAX_AGENTS>, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
      CUDA_TRY(cudaFuncSetAttribute((mg_obs_kernel<4, false>), cudaFuncAttributePreferredSharedMemoryCarveout, pct));
      CUDA_TRY(cudaFuncSetAttribute((mg_obs_kernel<4, false, false, true>), cudaFuncAttributePreferredSharedMemoryCarveout, pct));
      CUDA_TRY(cudaFuncSetAttribute((mg_obs_kernel<4, true>), cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    }
  }
#endif
  // slices of about 1024 envs: measured on B200, 4096 arena envs in the bench's steady state, 2 slices 935 M
  // agent-steps/s, 3: 934, 4: 954, 6: 945, 8: 921
  e->chunks = (num_envs + 1023) / 1024;
  if (e->chunks > 12) e->chunks = 12;
  if (const char* env_chunks = getenv("MG_CHUNKS")) e->chunks = atoi(env_chunks);  // tuning knob
  if (e->chunks < 1) e->chunks = 1;
  if (e->chunks > MG_MAX_CHUNKS) e->chunks = MG_MAX_CHUNKS;
  // with the host export in the graph (mg_step_host, export mode 1): measured 4 slices 185 M agent-steps/s, 12 slices 207-211
  e->chunks_host = (num_envs + 341) / 342;
  if (e->chunks_host > 12) e->chunks_host = 12;
  if (const char* env_chunks = getenv("MG_CHUNKS_HOST")) e->chunks_host = atoi(env_chunks);
  if (e->chunks_host < e->chunks) e->chunks_host = e->chunks;
  if (e->chunks_host > MG_MAX_CHUNKS) e->chunks_host = MG_MAX_CHUNKS;
  e->step_mode = MG_DEFAULT_STEP_MODE;
  if (const char* env_mode = getenv("MG_STEP_MODE"))  // tuning knob: "sliced", "queue", "fused"
    e->step_mode = strcmp(env_mode, "queue") == 0 ? 1 : (strcmp(env_mode, "fused") == 0 ? 2 : (strcmp(env_mode, "flags") == 0 ? 3 : 0));
  e->q_early = MG_QUEUE_OBS_CTAS_EARLY;
  e->q_late = MG_QUEUE_OBS_CTAS_LATE;
  if (const char* v = getenv("MG_QUEUE_EARLY")) e->q_early = atoi(v);
  if (const char* v = getenv("MG_QUEUE_LATE")) e->q_late = atoi(v);
#ifndef MG_SIMT_EMU
  CUDA_TRY(cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, device));
#else
  e->sm_count = 1;
#endif
  TRY(dev_alloc(e, &b.qctl, 1));
  TRY(dev_alloc(e, &b.queue, N));
  CUDA_TRY(cudaStreamCreateWithFlags(&e->capture_stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
  for (int i = 0; i < (e->chunks_host < 2 ? 2 : e->chunks_host); i++) {
    CUDA_TRY(cudaStreamCreateWithFlags(&e->chunk_stream[i], cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&e->ev_join[i], cudaEventDisableTiming));
  }
#ifndef MG_SIMT_EMU
  {
    int least = 0, greatest = 0;
    CUDA_TRY(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CUDA_TRY(cudaStreamCreateWithPriority(&e->prio_hi, cudaStreamNonBlocking, greatest));
    CUDA_TRY(cudaStreamCreateWithPriority(&e->prio_lo, cudaStreamNonBlocking, least));
  }
#endif
  *out = e;
  return MG_OK;
}

void mg_destroy(void* h) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return;
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  for (void* p : e->allocs) cudaFree(p);
  if (e->h_pinned) cudaFreeHost(e->h_pinned);
  if (e->h_export_bytes) cudaFreeHost(e->h_export_bytes);
  if (e->stream) cudaStreamDestroy(e->stream);
  if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
  if (e->graph) cudaGraphDestroy(e->graph);
  if (e->hgraph_exec) cudaGraphExecDestroy(e->hgraph_exec);
  if (e->hgraph) cudaGraphDestroy(e->hgraph);
  if (e->capture_stream) cudaStreamDestroy(e->capture_stream);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->prio_hi) cudaStreamDestroy(e->prio_hi);
  if (e->prio_lo) cudaStreamDestroy(e->prio_lo);
  for (int i = 0; i < MG_MAX_CHUNKS; i++) {
    if (e->ev_join[i]) cudaEventDestroy(e->ev_join[i]);
    if (e->chunk_stream[i]) cudaStreamDestroy(e->chunk_stream[i]);
  }
  delete e;
}

// Map pool, already split per kind by metta_b200/maps.py (host pointers):
// grids u16 [M][H*W], agents i32 [M][A][3], n_asm i32 [M], asms i32 [M][max_asm][3], n_chest, chests.
int mg_load_maps(void* h, int32_t num_maps, const uint16_t* grids, const int32_t* agents, const int32_t* n_asm,
                 const int32_t* asms, const int32_t* n_chest, const int32_t* chests) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || num_maps <= 0 || !grids || !agents || !n_asm || !asms || !n_chest || !chests)
    return fail(MG_ERR_ARG, "mg_load_maps: bad arguments");
  CUDA_TRY(cudaSetDevice(e->device));
  MgBatch& b = e->b;
  for (int m = 0; m < num_maps; m++)
    if (n_asm[m] > b.max_asm || n_chest[m] > b.max_chest || n_asm[m] < 0 || n_chest[m] < 0)
      return fail(MG_ERR_ARG, "mg_load_maps: a map has more assemblers/chests than the config's capacity");
  const size_t cells = (size_t)b.H * b.W, M = (size_t)num_maps;
  std::vector<uint16_t> padded(M * (size_t)b.grid_stride, 0);
  for (size_t m = 0; m < M; m++) memcpy(&padded[m * b.grid_stride], grids + m * cells, cells * sizeof(uint16_t));
  uint16_t* d_grid; int32_t *d_ag, *d_na, *d_as, *d_nc, *d_ch;
  int rc;
  TRY(dev_alloc(e, &d_grid, padded.size(), false));
  TRY(dev_alloc(e, &d_ag, M * b.A * 3, false));
  TRY(dev_alloc(e, &d_na, M, false));
  TRY(dev_alloc(e, &d_as, M * b.max_asm * 3, false));
  TRY(dev_alloc(e, &d_nc, M, false));
  TRY(dev_alloc(e, &d_ch, M * b.max_chest * 3, false));
  CUDA_TRY(cudaMemcpy(d_grid, padded.data(), padded.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(d_ag, agents, M * b.A * 3 * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(d_na, n_asm, M * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(d_as, asms, M * b.max_asm * 3 * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(d_nc, n_chest, M * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(d_ch, chests, M * b.max_chest * 3 * sizeof(int32_t), cudaMemcpyHostToDevice));
  {  // objects.<cell name>: walls, assemblers and chests per cell code and map (mettagrid_c.cpp:203,224,235)
    std::vector<float> counts(M * MG_MAX_CELLTYPES, 0.0f);
    for (size_t m = 0; m < M; m++) {
      for (size_t c = 0; c < cells; c++) {
        const uint16_t g = grids[m * cells + c];
        if (g != 0 && MG_GRID_KIND(g) == MG_GK_STATIC && g < MG_MAX_CELLTYPES) counts[m * MG_MAX_CELLTYPES + g] += 1.0f;
      }
      for (int i = 0; i < n_asm[m]; i++) {
        const int code = asms[(m * b.max_asm + i) * 3 + 2];
        if (code > 0 && code < MG_MAX_CELLTYPES) counts[m * MG_MAX_CELLTYPES + code] += 1.0f;
      }
      for (int i = 0; i < n_chest[m]; i++) {
        const int code = chests[(m * b.max_chest + i) * 3 + 2];
        if (code > 0 && code < MG_MAX_CELLTYPES) counts[m * MG_MAX_CELLTYPES + code] += 1.0f;
      }
    }
    float* d_counts;
    TRY(dev_alloc(e, &d_counts, counts.size(), false));
    CUDA_TRY(cudaMemcpy(d_counts, counts.data(), counts.size() * sizeof(float), cudaMemcpyHostToDevice));
    b.pool_obj_counts = d_counts;
  }
  {  // first-observation token counts per map: computed by the first env that starts an episode on the map
    uint16_t* d_t0; uint32_t* d_t0v;
    TRY(dev_alloc(e, &d_t0, M * b.A));
    TRY(dev_alloc(e, &d_t0v, M));
    b.pool_tok0 = d_t0; b.pool_tok0_valid = d_t0v;
  }
  b.pool_grid = d_grid; b.pool_agents = d_ag; b.pool_n_asm = d_na; b.pool_asm = d_as; b.pool_n_chest = d_nc; b.pool_chest = d_ch;
  b.num_maps = num_maps;
  e->maps_loaded = true;
  mark_graphs_dirty(e);
  return MG_OK;
}

// Bind caller-owned DEVICE buffers: obs u8 [N*A][T][3], terminals u8/bool [N*A], truncations, rewards f32, actions i32.
// Passing all nulls makes the library allocate them (used by the host-buffer path).
int mg_set_buffers(void* h, void* obs, void* terminals, void* truncations, void* rewards, void* actions) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_set_buffers: null handle");
  CUDA_TRY(cudaSetDevice(e->device));
  MgBatch& b = e->b;
  const size_t NA = (size_t)b.num_envs * b.A;
  if (!obs && !terminals && !truncations && !rewards && !actions) {
    if (!e->owns_buffers) {  // allocated once; every later all-NULL call binds the same buffers again
      int rc;
      TRY(dev_alloc(e, &e->own_obs, NA * b.T * 3 + 16));
      TRY(dev_alloc(e, &e->own_term, NA));
      TRY(dev_alloc(e, &e->own_trunc, NA));
      TRY(dev_alloc(e, &e->own_rew, NA));
      TRY(dev_alloc(e, &e->own_act, NA));
      e->owns_buffers = true;
    }
    b.obs = e->own_obs; b.terminals = e->own_term; b.truncations = e->own_trunc;
    b.rewards = e->own_rew; b.actions = e->own_act;
  } else {
    if (!obs || !terminals || !truncations || !rewards || !actions)
      return fail(MG_ERR_ARG, "mg_set_buffers: all five buffers are required");
    b.obs = (uint8_t*)obs; b.terminals = (uint8_t*)terminals; b.truncations = (uint8_t*)truncations;
    b.rewards = (float*)rewards; b.actions = (const int32_t*)actions;
  }
  e->buffers_bound = true;
  mark_graphs_dirty(e);
  return MG_OK;
}

static int launch_obs_slice(MgEnvs* e, const MgBatch& b, cudaStream_t s, const uint8_t* mask, bool count = true, bool wait_flags = false) {
  const dim3 grid((unsigned)(b.env_end - b.env_begin), (unsigned)((b.A + MG_OBS_WARPS - 1) / MG_OBS_WARPS));
  const int slot = (b.T * 3 + 16 + 15) & ~15;
  const int cpl = (b.obs_cells + 31) / 32;  // window cells per lane: 11x11 -> 4
  const int cpl_t = cpl <= 1 ? 1 : cpl <= 2 ? 2 : cpl <= 4 ? 4 : 8;
  const size_t smem = (size_t)(slot + MG_OBS_DESC_BYTES(cpl_t)) * MG_OBS_WARPS;
  if (wait_flags) {  // flag mode (arena-sized windows only; enqueue_flag_step checks)
    const dim3 wgrid(grid.y, grid.x);  // groups fastest: the CTAs of one env are dispatched together, envs in the order they finish
    MG_LAUNCH((mg_obs_kernel<4, false, false, true>), wgrid, MG_OBS_WARPS * 32, smem, s, b, mask);
  } else if (mask == nullptr && cpl_t == 4 && b.A < MG_OBS_WARPS) {  // few agents per env: rows back to back, no idle warps
    const dim3 fgrid((unsigned)(((b.env_end - b.env_begin) * b.A + MG_OBS_WARPS - 1) / MG_OBS_WARPS));
    MG_LAUNCH((mg_obs_kernel<4, false, true>), fgrid, MG_OBS_WARPS * 32, smem, s, b, mask);
  } else if (mask == nullptr) {
    if (cpl_t == 1) { MG_LAUNCH((mg_obs_kernel<1, false>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
    else if (cpl_t == 2) { MG_LAUNCH((mg_obs_kernel<2, false>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
    else if (cpl_t == 4) {
      // instantiations without the optional side outputs when none is bound, and for the reference's default row
      // of 200 tokens in an aligned buffer (sizes and loop bounds as constants)
      const bool aux = b.row_len != nullptr || b.max_flip != nullptr || b.obs_mask != nullptr;
      const bool t200 = b.T == 200 && (reinterpret_cast<uintptr_t>(b.obs) & 7) == 0;
      if (!aux && t200) { MG_LAUNCH((mg_obs_kernel<4, false, false, false, false, 200>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
      else if (!aux) { MG_LAUNCH((mg_obs_kernel<4, false, false, false, false>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
      else if (t200) { MG_LAUNCH((mg_obs_kernel<4, false, false, false, true, 200>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
      else { MG_LAUNCH((mg_obs_kernel<4, false>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
    }
    else { MG_LAUNCH((mg_obs_kernel<8, false>), grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
  } else if (cpl_t == 1) { MG_LAUNCH(mg_obs_kernel<1>, grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
  else if (cpl_t == 2) { MG_LAUNCH(mg_obs_kernel<2>, grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
  else if (cpl_t == 4) { MG_LAUNCH(mg_obs_kernel<4>, grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
  else { MG_LAUNCH(mg_obs_kernel<8>, grid, MG_OBS_WARPS * 32, smem, s, b, mask); }
  if (count) e->launches++;
  CUDA_TRY(cudaGetLastError());
  return MG_OK;
}

// The running maximum of the row lengths starts at zero for every full observation pass (a masked pass
// keeps the old value: it stays an upper bound).
static int begin_obs_pass(MgEnvs* e, cudaStream_t s) {
  if (e->b.max_flip != nullptr) CUDA_TRY(cudaMemsetAsync(e->b.max_flip, 0, sizeof(int32_t), s));
  return MG_OK;
}

static int launch_obs(MgEnvs* e, cudaStream_t s, const uint8_t* mask, bool after_step = false) {
  if (mask == nullptr) {
    const int rc = begin_obs_pass(e, s);
    if (rc != MG_OK) return rc;
  }
  if (!after_step) return launch_obs_slice(e, e->b, s, mask);
  MgBatch b = e->b;
  b.after_step = 1;
  return launch_obs_slice(e, b, s, mask);
}

static int launch_step_slice(MgEnvs* e, const MgBatch& b, cudaStream_t s) {
  const int grid = (b.env_end - b.env_begin + MG_STEP_WARPS - 1) / MG_STEP_WARPS;
  int cfg_bytes = ((b.cfg_words + 3) / 4) * 16;
  if (cfg_bytes > MG_CFG_SMEM_LIMIT) cfg_bytes = 0;
  if (b.A <= 32) {
    MG_LAUNCH(mg_step_kernel<32>, grid, MG_STEP_WARPS * 32, sizeof(StepSmem<32>) * MG_STEP_WARPS + (size_t)cfg_bytes, s, b, cfg_bytes);
  } else {
    MG_LAUNCH(mg_step_kernel<MG_MAX_AGENTS>, grid, MG_STEP_WARPS * 32, sizeof(StepSmem<MG_MAX_AGENTS>) * MG_STEP_WARPS + (size_t)cfg_bytes, s, b, cfg_bytes);
  }
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return MG_OK;
}

static int launch_export(MgEnvs* e, const MgBatch& b, cudaStream_t s, const MgHostTargets& ht, int row_begin, int row_end) {
  const int per_cta = MG_EXPORT_WARPS * MG_EXPORT_ROWS;
  const int grid = (row_end - row_begin + per_cta - 1) / per_cta;
  if (grid <= 0) return MG_OK;
  // the 64-byte comparison needs the three images congruent modulo 64
  uint8_t* shadow = e->d_shadow;
  if (ht.obs == nullptr || ((reinterpret_cast<uintptr_t>(b.obs) | reinterpret_cast<uintptr_t>(ht.obs) | reinterpret_cast<uintptr_t>(shadow)) & 63) != 0)
    shadow = nullptr;
  MG_LAUNCH(mg_export_rows_kernel, grid, MG_EXPORT_WARPS * 32, 0, s, (const uint8_t*)b.obs, (const float*)b.rewards, (const uint8_t*)b.terminals,
            (const uint8_t*)b.truncations, ht, (const uint16_t*)b.tok_counts, e->d_prev_counts, shadow, row_begin, row_end, b.T * 3,
            e->d_export_bytes);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return MG_OK;
}

// Fork from `s` to the chunk streams, run step + observation kernel (+ host export, if `ht`) of each
// env slice, join back.
static int enqueue_chunked_step(MgEnvs* e, cudaStream_t s, const MgHostTargets* ht = nullptr) {
  const int chunks = ht != nullptr ? e->chunks_host : e->chunks;
  const int n = e->b.num_envs, per = (n + chunks - 1) / chunks;
  int rc;
  if ((rc = begin_obs_pass(e, s)) != MG_OK) return rc;
  CUDA_TRY(cudaEventRecord(e->ev_fork, s));
  for (int c = 0; c < chunks; c++) {
    MgBatch slice = e->b;
    slice.env_begin = c * per;
    slice.env_end = (c + 1) * per < n ? (c + 1) * per : n;
    if (slice.env_begin >= slice.env_end) continue;
    cudaStream_t cs = e->chunk_stream[c];
    CUDA_TRY(cudaStreamWaitEvent(cs, e->ev_fork, 0));
    if ((rc = launch_step_slice(e, slice, cs)) != MG_OK) return rc;
    slice.after_step = 1;
    if ((rc = launch_obs_slice(e, slice, cs, nullptr)) != MG_OK) return rc;
    if (ht != nullptr && (rc = launch_export(e, slice, cs, *ht, slice.env_begin * slice.A, slice.env_end * slice.A)) != MG_OK) return rc;
    CUDA_TRY(cudaEventRecord(e->ev_join[c], cs));
    CUDA_TRY(cudaStreamWaitEvent(s, e->ev_join[c], 0));
  }
  return MG_OK;
}

#ifndef MG_SIMT_EMU
static int launch_obs_queue(MgEnvs* e, const MgBatch& b, cudaStream_t s, int ctas_per_sm) {
  const int slot = (b.T * 3 + 16 + 15) & ~15;
  const int cpl = (b.obs_cells + 31) / 32;
  const int cpl_t = cpl <= 1 ? 1 : cpl <= 2 ? 2 : cpl <= 4 ? 4 : 8;
  const size_t smem = (size_t)(slot + MG_OBS_DESC_BYTES(cpl_t)) * MG_OBS_WARPS;
  const int grid = e->sm_count * ctas_per_sm;
  if (cpl_t == 1) { MG_LAUNCH(mg_obs_queue_kernel<1>, grid, MG_OBS_WARPS * 32, smem, s, b); }
  else if (cpl_t == 2) { MG_LAUNCH(mg_obs_queue_kernel<2>, grid, MG_OBS_WARPS * 32, smem, s, b); }
  else if (cpl_t == 4) { MG_LAUNCH(mg_obs_queue_kernel<4>, grid, MG_OBS_WARPS * 32, smem, s, b); }
  else { MG_LAUNCH(mg_obs_queue_kernel<8>, grid, MG_OBS_WARPS * 32, smem, s, b); }
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return MG_OK;
}

// Queue mode: [begin] -> { step kernel (one env per CTA) -> late observation CTAs } beside { early
// observation CTAs }, joined back into `s`. Observation tasks are handed out in the order envs finish.
static int enqueue_queue_step(MgEnvs* e, cudaStream_t s) {
  MgBatch b = e->b;
  b.after_step = 1;
  int rc;
  if ((rc = begin_obs_pass(e, s)) != MG_OK) return rc;
  MG_LAUNCH(mg_queue_begin_kernel, 1, 32, 0, s, b.qctl);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaEventRecord(e->ev_fork, s));
  cudaStream_t sa = e->chunk_stream[0], sb = e->chunk_stream[1];
  CUDA_TRY(cudaStreamWaitEvent(sa, e->ev_fork, 0));
  CUDA_TRY(cudaStreamWaitEvent(sb, e->ev_fork, 0));
  const int n = b.env_end - b.env_begin;
  if (b.A <= 32) { MG_LAUNCH(mg_step_queue_kernel<32>, n, 32, sizeof(StepSmem<32>), sa, b); }
  else { MG_LAUNCH(mg_step_queue_kernel<MG_MAX_AGENTS>, n, 32, sizeof(StepSmem<MG_MAX_AGENTS>), sa, b); }
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  if (e->q_early > 0 && (rc = launch_obs_queue(e, b, sb, e->q_early)) != MG_OK) return rc;
  if (e->q_late > 0 && (rc = launch_obs_queue(e, b, sa, e->q_late)) != MG_OK) return rc;
  CUDA_TRY(cudaEventRecord(e->ev_join[0], sa));
  CUDA_TRY(cudaEventRecord(e->ev_join[1], sb));
  CUDA_TRY(cudaStreamWaitEvent(s, e->ev_join[0], 0));
  CUDA_TRY(cudaStreamWaitEvent(s, e->ev_join[1], 0));
  return MG_OK;
}
#endif

#ifndef MG_SIMT_EMU
// Flag mode: [begin] -> { step kernel, whole batch, on the high-priority stream } beside { observation kernel,
// whole batch, on the low-priority stream: every CTA waits for its env's mark }, joined back into `s`. The
// observation kernel keeps its shape (one short-lived warp per row); rows are encoded as soon as their env is
// done and an SM has room, instead of when the slowest env of a slice is done.
static bool flags_supported(const MgBatch& b) {
  const int cpl = (b.obs_cells + 31) / 32;
  return ((size_t)b.A * sizeof(AgentRec)) % 128 == 0 && cpl > 2 && cpl <= 4;
}
static int enqueue_flag_step(MgEnvs* e, cudaStream_t s) {
  MgBatch b = e->b;
  int rc;
  if ((rc = begin_obs_pass(e, s)) != MG_OK) return rc;
  MG_LAUNCH(mg_queue_begin_kernel, 1, 32, 0, s, b.qctl);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaEventRecord(e->ev_fork, s));
  CUDA_TRY(cudaStreamWaitEvent(e->prio_hi, e->ev_fork, 0));
  CUDA_TRY(cudaStreamWaitEvent(e->prio_lo, e->ev_fork, 0));
  b.env_flags = 1;
  if ((rc = launch_step_slice(e, b, e->prio_hi)) != MG_OK) return rc;
  b.after_step = 1;
  if ((rc = launch_obs_slice(e, b, e->prio_lo, nullptr, true, true)) != MG_OK) return rc;
  CUDA_TRY(cudaEventRecord(e->ev_join[0], e->prio_hi));
  CUDA_TRY(cudaEventRecord(e->ev_join[1], e->prio_lo));
  CUDA_TRY(cudaStreamWaitEvent(s, e->ev_join[0], 0));
  CUDA_TRY(cudaStreamWaitEvent(s, e->ev_join[1], 0));
  return MG_OK;
}
#endif

#ifndef MG_SIMT_EMU
}  // extern "C" (templates below)
// Fused mode: one persistent kernel steps the envs and encodes their observations (mg_fused_kernel).
template <int MAXA, int CPL>
static int launch_fused_t(MgEnvs* e, const MgBatch& b, cudaStream_t s) {
  const int slot = (b.T * 3 + 16 + 15) & ~15;
  size_t warp_bytes = sizeof(StepSmem<MAXA>);
  if ((size_t)(slot + MG_OBS_DESC_BYTES(CPL)) > warp_bytes) warp_bytes = (size_t)(slot + MG_OBS_DESC_BYTES(CPL));
  warp_bytes = (warp_bytes + 127) & ~(size_t)127;
  const size_t smem = warp_bytes * MG_FUSED_WARPS;
  static bool configured = false;  // per instantiation
  static int per_sm = 0;
  if (!configured) {
    CUDA_TRY(cudaFuncSetAttribute(mg_fused_kernel<MAXA, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = true;
  }
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mg_fused_kernel<MAXA, CPL>, MG_FUSED_WARPS * 32, smem));
  if (per_sm < 1) return fail(MG_ERR_CUDA, "fused step kernel does not fit an SM");
  const int n = b.env_end - b.env_begin;
  const int tasks = n * ((b.A + MG_FUSED_ROWS - 1) / MG_FUSED_ROWS);
  int grid = e->sm_count * per_sm;
  const int useful = ((n > tasks ? n : tasks) + MG_FUSED_WARPS - 1) / MG_FUSED_WARPS;
  if (grid > useful) grid = useful;
  MG_LAUNCH((mg_fused_kernel<MAXA, CPL>), grid, MG_FUSED_WARPS * 32, smem, s, b, (int)warp_bytes);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return MG_OK;
}

// The modes that hand an env to observation warps while other envs are still stepping rely on every cache
// line a consumer reads belonging to one env. Agent and global-token records are 64 bytes per agent,
// contiguous per env: whole 128-byte lines only for an even agent count. Otherwise the sliced graph is used.
static bool handover_supported(const MgBatch& b) { return ((size_t)b.A * sizeof(AgentRec)) % 128 == 0; }

static bool fused_supported(const MgBatch& b) {
  const int cpl = (b.obs_cells + 31) / 32;
  return handover_supported(b) && cpl <= 4;
}

static int enqueue_fused_step(MgEnvs* e, cudaStream_t s, const MgHostTargets* ht) {
  MgBatch b = e->b;
  b.after_step = 1;
  int rc;
  if ((rc = begin_obs_pass(e, s)) != MG_OK) return rc;
  MG_LAUNCH(mg_queue_begin_kernel, 1, 32, 0, s, b.qctl);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  const int cpl = (b.obs_cells + 31) / 32;
  if (b.A <= 32) rc = cpl <= 2 ? launch_fused_t<32, 2>(e, b, s) : launch_fused_t<32, 4>(e, b, s);
  else rc = launch_fused_t<MG_MAX_AGENTS, 4>(e, b, s);
  if (rc != MG_OK) return rc;
  if (ht != nullptr && (rc = launch_export(e, b, s, *ht, b.env_begin * b.A, b.env_end * b.A)) != MG_OK) return rc;
  return MG_OK;
}
extern "C" {
#endif

// Start episodes: seeds u32 [N], map index i32 [N], optional mask u8 [N] (host pointers; null mask = all).
// Writes the initial observation and clears rewards/terminals/truncations of the reset envs
// (MettaGrid ctor + set_buffers/_init_buffers, mettagrid_c.cpp:46-155,264-289).
int mg_reset(void* h, const uint32_t* seeds, const int32_t* map_idx, const uint8_t* mask) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !seeds || !map_idx) return fail(MG_ERR_ARG, "mg_reset: bad arguments");
  if (!e->maps_loaded) return fail(MG_ERR_STATE, "mg_reset: no maps loaded");
  if (!e->buffers_bound) return fail(MG_ERR_STATE, "mg_reset: buffers not bound (mg_set_buffers)");
  CUDA_TRY(cudaSetDevice(e->device));
  MgBatch& b = e->b;
  const size_t N = (size_t)b.num_envs;
  for (size_t i = 0; i < N; i++) {
    if (mask && !mask[i]) continue;
    if (map_idx[i] < 0 || map_idx[i] >= b.num_maps) return fail(MG_ERR_ARG, "mg_reset: map index out of range");
  }
  // A step may still be in flight on a non-blocking stream (mg_step is asynchronous): wait for it, then
  // hand seeds and map indices to the reset kernel, which patches the records of the envs it resets on the
  // device. Records of unmasked envs are never rewritten from the host.
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaMemcpy(e->d_seeds, seeds, N * sizeof(uint32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(e->d_map_idx, map_idx, N * sizeof(int32_t), cudaMemcpyHostToDevice));
  const uint8_t* d_mask = nullptr;
  if (mask) {
    CUDA_TRY(cudaMemcpy(e->d_mask, mask, N, cudaMemcpyHostToDevice));
    d_mask = e->d_mask;
  }
  const int grid = (b.num_envs + MG_STEP_WARPS - 1) / MG_STEP_WARPS;
  MG_LAUNCH(mg_reset_kernel, grid, MG_STEP_WARPS * 32, 0, (cudaStream_t)0, b, d_mask, (const uint32_t*)e->d_seeds, (const int32_t*)e->d_map_idx);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  int rc;
  TRY(launch_obs(e, (cudaStream_t)0, d_mask));
  CUDA_TRY(cudaDeviceSynchronize());
  return MG_OK;
}

// One step of every env on `stream` (a cudaStream_t, may be null): reads the bound actions,
// overwrites observations/rewards/terminals/truncations. Asynchronous.
#ifndef MG_SIMT_EMU
static int capture_step_graph(MgEnvs* e, const MgHostTargets* ht, cudaGraph_t* graph, cudaGraphExec_t* exec) {
  if (*exec) { cudaGraphExecDestroy(*exec); *exec = nullptr; }
  if (*graph) { cudaGraphDestroy(*graph); *graph = nullptr; }
  const long long before = e->launches;
  CUDA_TRY(cudaStreamBeginCapture(e->capture_stream, cudaStreamCaptureModeThreadLocal));
  int rc;
  if (e->step_mode == 3 && ht == nullptr && flags_supported(e->b)) rc = enqueue_flag_step(e, e->capture_stream);
  else if (e->step_mode == 2 && fused_supported(e->b)) rc = enqueue_fused_step(e, e->capture_stream, ht);
  else if (e->step_mode == 1 && ht == nullptr && handover_supported(e->b)) rc = enqueue_queue_step(e, e->capture_stream);
  else rc = enqueue_chunked_step(e, e->capture_stream, ht);
  cudaError_t end = cudaStreamEndCapture(e->capture_stream, graph);
  (ht != nullptr ? e->hgraph_launches : e->graph_launches) = e->launches - before;  // kernels one replay of this graph launches
  e->launches = before;
  if (rc != MG_OK) return rc;
  if (end != cudaSuccess) return fail(MG_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(end));
  CUDA_TRY(cudaGraphInstantiate(exec, *graph, 0));
  return MG_OK;
}
#endif

int mg_step(void* h, void* stream) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_step: null handle");
  if (!e->maps_loaded || !e->buffers_bound) return fail(MG_ERR_STATE, "mg_step: maps or buffers missing");
  CUDA_TRY(cudaSetDevice(e->device));
  cudaStream_t s = (cudaStream_t)stream;
  if (e->chunks <= 1 && e->step_mode == 0) {
    int rc = launch_step_slice(e, e->b, s);
    return rc != MG_OK ? rc : launch_obs(e, s, nullptr, true);
  }
#ifdef MG_SIMT_EMU
  return enqueue_chunked_step(e, s);
#else
  if (e->graph_dirty) {  // (re)capture: buffers, options or maps changed since the last capture
    int rc = capture_step_graph(e, nullptr, &e->graph, &e->graph_exec);
    if (rc != MG_OK) return rc;
    e->graph_dirty = false;
  }
  CUDA_TRY(cudaGraphLaunch(e->graph_exec, s));
  e->launches += e->graph_launches;
  return MG_OK;
#endif
}

// One synchronous step with CUDA events around each kernel, for roofline measurements:
// out_ms[0] = action/step kernel, out_ms[1] = observation kernel (milliseconds).
int mg_profile_step(void* h, void* stream, float* out_ms) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !out_ms) return fail(MG_ERR_ARG, "mg_profile_step: bad arguments");
  if (!e->maps_loaded || !e->buffers_bound) return fail(MG_ERR_STATE, "mg_profile_step: maps or buffers missing");
  CUDA_TRY(cudaSetDevice(e->device));
  MgBatch& b = e->b;
  cudaStream_t s = (cudaStream_t)stream;
  cudaEvent_t ev[3];
  for (int i = 0; i < 3; i++) CUDA_TRY(cudaEventCreate(&ev[i]));
  int rc;
  CUDA_TRY(cudaEventRecord(ev[0], s));
  TRY(launch_step_slice(e, b, s));
  CUDA_TRY(cudaEventRecord(ev[1], s));
  TRY(launch_obs(e, s, nullptr, true));
  CUDA_TRY(cudaEventRecord(ev[2], s));
  CUDA_TRY(cudaEventSynchronize(ev[2]));
  CUDA_TRY(cudaEventElapsedTime(&out_ms[0], ev[0], ev[1]));
  CUDA_TRY(cudaEventElapsedTime(&out_ms[1], ev[1], ev[2]));
  for (int i = 0; i < 3; i++) cudaEventDestroy(ev[i]);
  return MG_OK;
}

// Recompute observations only (debugging / after mg_write).
int mg_compute_observations(void* h, void* stream) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !e->buffers_bound) return fail(MG_ERR_STATE, "mg_compute_observations: buffers missing");
  CUDA_TRY(cudaSetDevice(e->device));
  return launch_obs(e, (cudaStream_t)stream, nullptr);
}

// Is this host pointer page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory)? Then the DMA
// engines can reach it directly and no staging copy is needed.
static bool is_pinned_host(const void* p) {
#ifdef MG_SIMT_EMU
  (void)p;
  return false;
#else
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeHost;
#endif
}

// Host-buffer step: the call a numpy caller makes. actions i32 [N*A] in; obs, rewards, terminals,
// truncations out. Page-locked caller buffers are used directly by the copy engines; pageable ones
// go through the library's pinned staging buffer. Runs on the library's stream; returns when done.
// mg_step_host in two halves: _begin enqueues the copies and the step on the handle's own stream and returns,
// _end waits for them. A caller that keeps two handles (two halves of its envs) in flight has one half's
// observations crossing the host link while the other half steps.
int mg_step_host_begin(void* h, const int32_t* actions, uint8_t* obs, float* rewards, uint8_t* terminals, uint8_t* truncations) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !actions || !obs || !rewards || !terminals || !truncations) return fail(MG_ERR_ARG, "mg_step_host: null pointer");
  if (!e->buffers_bound) return fail(MG_ERR_STATE, "mg_step_host: call mg_set_buffers first (all nulls = library-owned)");
  if (e->host_pending.active) return fail(MG_ERR_STATE, "mg_step_host_begin: the previous mg_step_host_begin has not been ended");
  CUDA_TRY(cudaSetDevice(e->device));
  MgBatch& b = e->b;
  const size_t NA = (size_t)b.num_envs * b.A;
  const size_t obs_bytes = NA * b.T * 3;
  const bool direct = is_pinned_host(actions) && is_pinned_host(obs) && is_pinned_host(rewards) && is_pinned_host(terminals) &&
                      is_pinned_host(truncations);
  const size_t off_act = 0, off_rew = off_act + NA * 4, off_term = off_rew + NA * 4, off_trunc = off_term + NA;
  const size_t off_obs = (off_trunc + NA + 255) & ~(size_t)255, total = off_obs + obs_bytes;
  if (!direct && e->h_pinned_bytes < total) {
    if (e->h_pinned) cudaFreeHost(e->h_pinned);
    CUDA_TRY(cudaMallocHost((void**)&e->h_pinned, total));
    e->h_pinned_bytes = total;
  }
  cudaStream_t s = e->stream;
  uint8_t* st = e->h_pinned;
  const void* src_act = actions;
  if (!direct) { memcpy(st + off_act, actions, NA * 4); src_act = st + off_act; }
  CUDA_TRY(cudaMemcpyAsync((void*)b.actions, src_act, NA * 4, cudaMemcpyHostToDevice, s));
  e->h2d_total += NA * 4;
  e->d2h_total += NA * 6;
  int rc;
  // observations: the whole buffer, or (export mode 1, same page-locked buffers as last time) only the
  // prefix of each row that can differ from what the host holds, written over PCIe by a kernel that
  // follows each slice's observation kernel inside the step graph (rewards and flags ride along)
  const int row_bytes = b.T * 3;
  MgHostTargets ht;
  ht.obs = nullptr; ht.rewards = nullptr; ht.terminals = nullptr; ht.truncations = nullptr;
  bool sparse = false;
#ifndef MG_SIMT_EMU
  if (direct && e->host_export_mode == 1 && e->host_obs_last == obs && row_bytes % 8 == 0 &&
      (reinterpret_cast<uintptr_t>(b.obs) & 7) == 0 && (reinterpret_cast<uintptr_t>(obs) & 7) == 0) {
    void *p0 = nullptr, *p1 = nullptr, *p2 = nullptr, *p3 = nullptr;
    if (cudaHostGetDevicePointer(&p0, obs, 0) == cudaSuccess && cudaHostGetDevicePointer(&p1, rewards, 0) == cudaSuccess &&
        cudaHostGetDevicePointer(&p2, terminals, 0) == cudaSuccess && cudaHostGetDevicePointer(&p3, truncations, 0) == cudaSuccess) {
      ht.obs = reinterpret_cast<uint8_t*>(p0); ht.rewards = reinterpret_cast<float*>(p1);
      ht.terminals = reinterpret_cast<uint8_t*>(p2); ht.truncations = reinterpret_cast<uint8_t*>(p3);
      sparse = true;
    } else {
      cudaGetLastError();
    }
  }
  if (sparse) {
    if (e->chunks_host <= 1 && e->step_mode == 0) {
      TRY(launch_step_slice(e, b, s));
      TRY(launch_obs(e, s, nullptr, true));
      TRY(launch_export(e, b, s, ht, 0, (int)NA));
    } else {
      if (e->hgraph_dirty || memcmp(&ht, &e->hgraph_targets, sizeof(ht)) != 0) {
        TRY(capture_step_graph(e, &ht, &e->hgraph, &e->hgraph_exec));
        e->hgraph_targets = ht;
        e->hgraph_dirty = false;
      }
      CUDA_TRY(cudaGraphLaunch(e->hgraph_exec, s));
      e->launches += e->hgraph_launches;
    }
    CUDA_TRY(cudaMemcpyAsync(e->h_export_bytes, e->d_export_bytes, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    e->host_pending = MgEnvs::HostPending();
    e->host_pending.active = true;
    return MG_OK;
  }
#endif
  TRY(mg_step(h, (void*)s));
  CUDA_TRY(cudaMemcpyAsync(direct ? (void*)obs : (void*)(st + off_obs), b.obs, obs_bytes, cudaMemcpyDeviceToHost, s));
  e->d2h_total += obs_bytes;
  e->host_obs_last = nullptr;
  if (direct && e->host_export_mode == 1) {  // from now on the host buffer is known row by row (and byte by byte)
    TRY(launch_export(e, b, s, ht, 0, (int)NA));
    if (e->d_shadow) CUDA_TRY(cudaMemcpyAsync(e->d_shadow, b.obs, obs_bytes, cudaMemcpyDeviceToDevice, s));
    e->host_obs_last = obs;
  }
  CUDA_TRY(cudaMemcpyAsync(direct ? (void*)rewards : (void*)(st + off_rew), b.rewards, NA * 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(direct ? (void*)terminals : (void*)(st + off_term), b.terminals, NA, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(direct ? (void*)truncations : (void*)(st + off_trunc), b.truncations, NA, cudaMemcpyDeviceToHost, s));
  MgEnvs::HostPending& hp = e->host_pending;
  hp = MgEnvs::HostPending();
  hp.active = true;
  hp.staged = !direct;
  hp.obs = obs; hp.rewards = rewards; hp.terminals = terminals; hp.truncations = truncations;
  hp.NA = NA; hp.obs_bytes = obs_bytes; hp.off_obs = off_obs; hp.off_rew = off_rew; hp.off_term = off_term; hp.off_trunc = off_trunc;
  return MG_OK;
}

int mg_step_host_end(void* h) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_step_host_end: null handle");
  if (!e->host_pending.active) return fail(MG_ERR_STATE, "mg_step_host_end: no mg_step_host_begin to end");
  CUDA_TRY(cudaSetDevice(e->device));
  const MgEnvs::HostPending hp = e->host_pending;
  e->host_pending.active = false;
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  if (hp.staged) {
    const uint8_t* st = e->h_pinned;
    memcpy(hp.obs, st + hp.off_obs, hp.obs_bytes);
    memcpy(hp.rewards, st + hp.off_rew, hp.NA * 4);
    memcpy(hp.terminals, st + hp.off_term, hp.NA);
    memcpy(hp.truncations, st + hp.off_trunc, hp.NA);
  }
  return check_queue_error(e, "mg_step_host");
}

int mg_step_host(void* h, const int32_t* actions, uint8_t* obs, float* rewards, uint8_t* terminals, uint8_t* truncations) {
  const int rc = mg_step_host_begin(h, actions, obs, rewards, terminals, truncations);
  if (rc != MG_OK) return rc;
  return mg_step_host_end(h);
}

// How mg_step_host delivers observations into a page-locked caller buffer. 0 (default): the whole
// buffer every call. 1: after one full copy, only the leading part of each row that can differ from
// what the buffer holds (rows end in 0xFF padding, and the padding does not move unless the token count
// does). Mode 1 relies on the library being the only writer of that buffer between calls; passing
// another buffer, or calling this again, starts over with a full copy.
int mg_set_host_export(void* h, int32_t mode) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || (mode != 0 && mode != 1)) return fail(MG_ERR_ARG, "mg_set_host_export: mode must be 0 or 1");
  e->host_export_mode = mode;
  e->host_obs_last = nullptr;
  if (mode == 1 && e->d_shadow == nullptr) {  // allocated here so that mg_step_host itself never allocates
    CUDA_TRY(cudaSetDevice(e->device));
    int rc;
    TRY(dev_alloc(e, &e->d_shadow, (size_t)e->b.num_envs * e->b.A * e->b.T * 3, false));
  }
  return MG_OK;
}

// Bytes mg_step_host moved over the host link since mg_create (what the copies and the export kernel wrote).
int mg_transfer_bytes(void* h, uint64_t* h2d, uint64_t* d2h) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !h2d || !d2h) return fail(MG_ERR_ARG, "mg_transfer_bytes: bad arguments");
  *h2d = e->h2d_total;
  *d2h = e->d2h_total + (e->h_export_bytes ? *e->h_export_bytes : 0ull);
  return MG_OK;
}

// Optional side outputs of every observation pass, in caller-owned device memory (any of them may be null):
// row_lengths i32 [N*A], max_flip i32 [1], obs_mask u8 [N*A][T]. See include/mettagrid_b200.h.
int mg_set_obs_aux(void* h, void* row_lengths, void* max_flip, void* obs_mask) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_set_obs_aux: null handle");
  if ((reinterpret_cast<uintptr_t>(row_lengths) | reinterpret_cast<uintptr_t>(max_flip)) & 3)
    return fail(MG_ERR_ARG, "mg_set_obs_aux: row_lengths and max_flip must be 4-byte aligned");
  if (reinterpret_cast<uintptr_t>(obs_mask) & 7) return fail(MG_ERR_ARG, "mg_set_obs_aux: obs_mask must be 8-byte aligned");
  e->b.row_len = reinterpret_cast<int32_t*>(row_lengths);
  e->b.max_flip = reinterpret_cast<int32_t*>(max_flip);
  e->b.obs_mask = reinterpret_cast<uint8_t*>(obs_mask);
  mark_graphs_dirty(e);
  return MG_OK;
}

// auto_reset: 1 = MettaGridPufferEnv semantics (default), 0 = bare MettaGrid (keeps stepping past the end).
// map_stride: on auto-reset the env moves to map (index + stride) mod pool size.
int mg_set_options(void* h, int32_t auto_reset, int32_t map_stride) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_set_options: null handle");
  e->b.auto_reset = auto_reset ? 1 : 0;
  e->b.map_stride = map_stride < 0 ? 0 : map_stride;
  mark_graphs_dirty(e);
  return MG_OK;
}

// EarlyResetHandler: truncate the running episode of env i when current_step >= steps[i] (0 = never).
int mg_set_early_reset(void* h, const uint32_t* steps) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !steps) return fail(MG_ERR_ARG, "mg_set_early_reset: bad arguments");
  CUDA_TRY(cudaSetDevice(e->device));
  const size_t N = (size_t)e->b.num_envs;
  std::vector<EnvRec> recs(N);
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaMemcpy(recs.data(), e->b.env, N * sizeof(EnvRec), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; i++) recs[i].early_reset_step = steps[i];
  CUDA_TRY(cudaMemcpy(e->b.env, recs.data(), N * sizeof(EnvRec), cudaMemcpyHostToDevice));
  return MG_OK;
}

// Raw state access for the conformance harness and infos. `what`:
enum { MG_ST_ENV = 0, MG_ST_GRID = 1, MG_ST_AGENTS = 2, MG_ST_ASMS = 3, MG_ST_CHESTS = 4, MG_ST_AGENT_STATS = 5,
       MG_ST_GAME_STATS = 6, MG_ST_LAST_ACTION = 7, MG_ST_ACTION_SUCCESS = 8, MG_ST_EPISODE_ACCUM = 9, MG_ST_RNG = 10,
       MG_ST_OBS = 11, MG_ST_REWARDS = 12, MG_ST_TERMINALS = 13, MG_ST_TRUNCATIONS = 14, MG_ST_ACTIONS = 15,
       MG_ST_DEBUG = 16, MG_ST_INVALID_IDS = 17, MG_ST_TOKEN_RECORDS = 18, MG_ST_COUNT = 19 };

static void* state_ptr(MgEnvs* e, int what, size_t* bytes) {
  MgBatch& b = e->b;
  const size_t N = (size_t)b.num_envs, A = (size_t)b.A;
  switch (what) {
    case MG_ST_ENV: *bytes = N * sizeof(EnvRec); return b.env;
    case MG_ST_GRID: *bytes = N * b.grid_stride * sizeof(uint16_t); return b.grid;
    case MG_ST_AGENTS: *bytes = N * A * sizeof(AgentRec); return b.agents;
    case MG_ST_ASMS: *bytes = N * b.max_asm * sizeof(AsmRec); return b.asms;
    case MG_ST_CHESTS: *bytes = N * b.max_chest * sizeof(ChestRec); return b.chests;
    case MG_ST_AGENT_STATS: *bytes = N * A * b.S * sizeof(float); return b.agent_stats;
    case MG_ST_GAME_STATS: *bytes = N * b.GS * sizeof(float); return b.game_stats;
    case MG_ST_LAST_ACTION: *bytes = N * A * sizeof(int32_t); return b.last_action;
    case MG_ST_ACTION_SUCCESS: *bytes = N * A; return b.action_success;
    case MG_ST_EPISODE_ACCUM: *bytes = (size_t)(2 + b.S + b.GS) * sizeof(double); return b.episode_accum;
    case MG_ST_RNG: *bytes = N * MT_N * sizeof(uint32_t); return b.rng;
    case MG_ST_OBS: *bytes = N * A * b.T * 3; return b.obs;
    case MG_ST_REWARDS: *bytes = N * A * sizeof(float); return b.rewards;
    case MG_ST_TERMINALS: *bytes = N * A; return b.terminals;
    case MG_ST_TRUNCATIONS: *bytes = N * A; return b.truncations;
    case MG_ST_ACTIONS: *bytes = N * A * sizeof(int32_t); return (void*)b.actions;
    case MG_ST_DEBUG: *bytes = b.dbg ? N * 16 * sizeof(uint32_t) : 0; return b.dbg;
    case MG_ST_TOKEN_RECORDS: *bytes = (N * (size_t)b.rec_slots) << b.rec_shift; return b.etok;
    case MG_ST_INVALID_IDS: *bytes = N * A * MG_INVALID_WORDS * sizeof(int32_t); return b.invalid_ids;
    default: *bytes = 0; return nullptr;
  }
}

int64_t mg_sizeof(void* h, int32_t what) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  size_t bytes = 0;
  if (!e || !state_ptr(e, what, &bytes)) return -1;
  return (int64_t)bytes;
}

int mg_read(void* h, int32_t what, void* host_out, int64_t nbytes) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !host_out) return fail(MG_ERR_ARG, "mg_read: bad arguments");
  CUDA_TRY(cudaSetDevice(e->device));
  size_t bytes = 0;
  void* p = state_ptr(e, what, &bytes);
  if (!p) return fail(MG_ERR_ARG, "mg_read: unknown or unbound state array");
  if ((size_t)nbytes != bytes) return fail(MG_ERR_ARG, "mg_read: size mismatch, expected " + std::to_string(bytes));
  CUDA_TRY(cudaDeviceSynchronize());
  {
    const int qrc = check_queue_error(e, "mg_read");
    if (qrc != MG_OK) return qrc;
  }
  if (what == MG_ST_GAME_STATS) {  // pending token counts of the last observation pass
    const int grid = (e->b.num_envs + MG_STEP_WARPS - 1) / MG_STEP_WARPS;
    MG_LAUNCH(mg_fold_kernel, grid, MG_STEP_WARPS * 32, 0, (cudaStream_t)0, e->b);
    e->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaDeviceSynchronize());
  }
  CUDA_TRY(cudaMemcpy(host_out, p, bytes, cudaMemcpyDeviceToHost));
  if (what == MG_ST_EPISODE_ACCUM) {  // kept as 48.16 fixed point on the device, handed out as doubles
    long long* raw = reinterpret_cast<long long*>(host_out);
    double* out = reinterpret_cast<double*>(host_out);
    for (size_t i = 0; i < bytes / sizeof(double); i++) out[i] = (double)raw[i] / 65536.0;
  }
  if (what == MG_ST_AGENT_STATS) {  // slots 0..MG_HOT_STATS-1 of every agent live in an array of their own
    const size_t NA = (size_t)e->b.num_envs * e->b.A;
    std::vector<float> hot(NA * MG_HOT_STATS);
    CUDA_TRY(cudaMemcpy(hot.data(), e->b.hot_stats, hot.size() * sizeof(float), cudaMemcpyDeviceToHost));
    float* out = reinterpret_cast<float*>(host_out);
    for (size_t a = 0; a < NA; a++) memcpy(out + a * e->b.S, hot.data() + a * MG_HOT_STATS, MG_HOT_STATS * sizeof(float));
  }
  return MG_OK;
}

int mg_write(void* h, int32_t what, const void* host_in, int64_t nbytes) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !host_in) return fail(MG_ERR_ARG, "mg_write: bad arguments");
  CUDA_TRY(cudaSetDevice(e->device));
  size_t bytes = 0;
  void* p = state_ptr(e, what, &bytes);
  if (!p) return fail(MG_ERR_ARG, "mg_write: unknown or unbound state array");
  if ((size_t)nbytes != bytes) return fail(MG_ERR_ARG, "mg_write: size mismatch, expected " + std::to_string(bytes));
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaMemcpy(p, host_in, bytes, cudaMemcpyHostToDevice));
  if (what == MG_ST_AGENT_STATS) {
    const size_t NA = (size_t)e->b.num_envs * e->b.A;
    std::vector<float> hot(NA * MG_HOT_STATS);
    const float* in = reinterpret_cast<const float*>(host_in);
    for (size_t a = 0; a < NA; a++) memcpy(hot.data() + a * MG_HOT_STATS, in + a * e->b.S, MG_HOT_STATS * sizeof(float));
    CUDA_TRY(cudaMemcpy(e->b.hot_stats, hot.data(), hot.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  return MG_OK;
}

// MettaGrid::set_inventory (mettagrid_c.cpp:959-965): items/amounts i32[n] in the caller's dict order.
int mg_set_inventory(void* h, int32_t env, int32_t agent, const int32_t* items_amounts /* [2][n] */, int32_t n) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || n < 0 || n > MG_RES_STRIDE || (n > 0 && !items_amounts)) return fail(MG_ERR_ARG, "mg_set_inventory: bad arguments");
  if (env < 0 || env >= e->b.num_envs) return fail(MG_ERR_ARG, "mg_set_inventory: env out of range");
  if (agent < 0 || agent >= e->b.A) return MG_OK;  // the reference ignores unknown agent ids
  for (int i = 0; i < n; i++)
    if (items_amounts[i] < 0 || items_amounts[i] >= e->b.R) return fail(MG_ERR_ARG, "mg_set_inventory: unknown resource id");
  CUDA_TRY(cudaSetDevice(e->device));
  CUDA_TRY(cudaDeviceSynchronize());
  if (n > 0) CUDA_TRY(cudaMemcpy(e->d_tmp, items_amounts, sizeof(int32_t) * 2 * (size_t)n, cudaMemcpyHostToDevice));
  MG_LAUNCH(mg_set_inventory_kernel, 1, 32, 0, (cudaStream_t)0, e->b, env, agent, e->d_tmp, e->d_tmp + n, n);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaDeviceSynchronize());
  return MG_OK;
}

// Layout of the token records (mg_read what=18): out[0] = bytes per record, out[1] = records per env, out[2..4] = first
// record of the agents, assemblers and chests of an env (records before out[2] belong to the map cell codes).
int mg_token_record_layout(void* h, int32_t* out5) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !out5) return fail(MG_ERR_ARG, "mg_token_record_layout: bad arguments");
  const MgBatch& b = e->b;
  out5[0] = 1 << b.rec_shift;
  out5[1] = b.rec_slots;
  out5[2] = (int32_t)(b.rec_kind_base_lo >> 16);
  out5[3] = (int32_t)(b.rec_kind_base_hi & 0xFFFFu);
  out5[4] = (int32_t)(b.rec_kind_base_hi >> 16);
  return MG_OK;
}

int64_t mg_kernel_launches(void* h) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  return e ? e->launches : -1;
}

int mg_sync(void* h) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e) return fail(MG_ERR_ARG, "mg_sync: null handle");
  CUDA_TRY(cudaSetDevice(e->device));
  CUDA_TRY(cudaDeviceSynchronize());
  return check_queue_error(e, "mg_sync");
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// mg_peek_observations_noalloc: observations of the current state into a caller's side buffer,
// without allocating anything and without disturbing the episode's token statistics.
//
// Who needs this: a monitor, evaluator or replay writer that wants to look at what the agents see
// in the middle of a rollout while the trainer owns the GPU. Such a caller often may not allocate
// device memory at all: the training framework's caching allocator has reserved nearly all of it
// (a foreign cudaMalloc fails or forces a device-wide synchronisation), or the caller runs in a
// thread whose latency budget forbids the implicit sync of cudaMalloc/cudaFree. The regular
// observation pass cannot simply be pointed at another output either, because it leaves per-agent
// token counts in scratch memory that the next step folds into tokens_written / tokens_dropped /
// tokens_free_space; a second pass would be counted twice.
//
// So this variant of the observation kernel keeps that scratch (two u16 counts per agent) in the
// file-scope device array below instead of cudaMalloc'ed memory. The array exists once per GPU
// that loads this library and stays until the process exits: MG_PEEK_MAX_AGENTS * 4 bytes = 1 MiB.
// Larger batches are refused. All calls on a device share the array and it keeps whatever the
// last call wrote; nothing in it has to start at zero, because every slot a call reads was written
// earlier in that same call. Two calls that overlap on the GPU (different streams, or different
// handles on one device) race on it: serialise them.
//
// The name `g_peek_scratch` is only a host-side stand-in in host code, not a device address. The
// first call on each device therefore asks the runtime for the real address
// (cudaGetSymbolAddress, with that device current) and remembers it; make that first call before
// starting any stream capture.
#define MG_PEEK_MAX_AGENTS 262144
#define MG_PEEK_MAX_DEVICES 64
static __device__ __align__(256) uint16_t g_peek_scratch[MG_PEEK_MAX_AGENTS * 2];
static uint16_t* g_peek_addr[MG_PEEK_MAX_DEVICES];  // device address of g_peek_scratch, per device ordinal

extern "C" int mg_peek_observations_noalloc(void* h, void* obs_out_dev, void* stream) {
  MgEnvs* e = reinterpret_cast<MgEnvs*>(h);
  if (!e || !obs_out_dev) return fail(MG_ERR_ARG, "mg_peek_observations_noalloc: null handle or output buffer");
  if (!e->maps_loaded) return fail(MG_ERR_STATE, "mg_peek_observations_noalloc: no maps loaded, nothing to observe");
  const MgBatch& b = e->b;
  const long long agents = (long long)b.num_envs * b.A;
  if (agents > MG_PEEK_MAX_AGENTS)
    return fail(MG_ERR_ARG, "mg_peek_observations_noalloc: " + std::to_string(agents) + " agents exceed the static scratch of " +
                                std::to_string(MG_PEEK_MAX_AGENTS) + " agents; use mg_compute_observations on a bound buffer");
  if (e->device < 0 || e->device >= MG_PEEK_MAX_DEVICES) return fail(MG_ERR_ARG, "mg_peek_observations_noalloc: device ordinal out of range");
  CUDA_TRY(cudaSetDevice(e->device));
  uint16_t* scratch = g_peek_addr[e->device];
  if (scratch == nullptr) {  // once per device, with that device current
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_peek_scratch));
    if (p == nullptr) return fail(MG_ERR_CUDA, "mg_peek_observations_noalloc: cudaGetSymbolAddress returned no address");
    scratch = g_peek_addr[e->device] = reinterpret_cast<uint16_t*>(p);
  }
  MgBatch view = b;              // same state, other destinations
  view.obs = reinterpret_cast<uint8_t*>(obs_out_dev);
  view.tok_counts = scratch;
  view.row_len = nullptr; view.max_flip = nullptr; view.obs_mask = nullptr;  // the bound side outputs stay as they are
  cudaStream_t s = (cudaStream_t)stream;
  const int rc_launch = launch_obs_slice(e, view, s, nullptr, false);  // the handle itself is not written to: no launch count
  if (rc_launch != MG_OK) return rc_launch;
  return MG_OK;
}
