/* mettagrid_b200.h -- C-ABI of libmettagrid_b200.so: a batched MettaGrid simulator for B200.
 *
 * This is the boundary a maintainer of the reference binds instead of the pybind11 class
 * `MettaGrid` (packages/mettagrid/cpp/bindings/mettagrid_c.cpp:1007-1041). One handle holds N
 * environments of one lowered game config (include/mg_config.h) on one GPU. All entry points
 * return 0 on success; otherwise mg_last_error() holds the message (the reference throws
 * std::runtime_error at the same places, mettagrid_c.cpp:65-70,191-193,680-719,743-749).
 * Pointers named *_dev are device memory owned by the caller; *_host are host memory.
 * Not thread-safe per handle, like the reference core (SURVEY.md 8b "threading").
 */
#ifndef METTAGRID_B200_H_
#define METTAGRID_B200_H_

#include <stdint.h>

#include "mg_config.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* mg_last_error(void);
int mg_abi_version(void); /* == MG_CONFIG_VERSION */

/* Host helpers used while lowering a config: iteration order of the libstdc++ unordered_map the
 * reference would hold for these keys (replaces nothing in the reference: it *is* the reference's
 * container, see metta_b200/csrc/mg_host_order.cpp; pybind11/stl.h map_caster::load reserves first). */
int mg_umap_order_u8(const uint8_t* keys, int n, int reserve_first, int32_t* out_positions);
int mg_umap_order_str(const char* keys_nul_separated, int n, int reserve_first, int32_t* out_positions);

/* MettaGrid::MettaGrid(GameConfig, map, seed) (mettagrid_c.cpp:46-155), for num_envs environments.
 * config: lowered blob; maps and seeds are supplied by mg_load_maps / mg_reset. */
int mg_create(const int32_t* config_words, int64_t n_words, int32_t num_envs, int32_t device, void** out_handle);
void mg_destroy(void* handle);

/* The `map` constructor argument (list[list[str]], mettagrid_c.cpp:159-245), as a pool of maps
 * pre-split per object kind (metta_b200/maps.py build_pool); kept in HBM for auto-reset
 * (replaces SharedMapCache, mettagrid/simulator/map_cache.py). */
int mg_load_maps(void* handle, int32_t num_maps, const uint16_t* grids_host, const int32_t* agents_host,
                 const int32_t* n_assemblers_host, const int32_t* assemblers_host, const int32_t* n_chests_host,
                 const int32_t* chests_host);

/* MettaGrid::set_buffers (mettagrid_c.cpp:721-738): bind caller-owned buffers for all N*A agents:
 * observations u8[N*A][T][3], terminals u8[N*A], truncations u8[N*A], rewards f32[N*A],
 * actions i32[N*A]. All NULL = let the library own them (for mg_step_host). */
int mg_set_buffers(void* handle, void* obs_dev, void* terminals_dev, void* truncations_dev, void* rewards_dev,
                   void* actions_dev);

/* New episodes: the constructor's seed argument and map choice per env, then the initial
 * observation of _init_buffers (mettagrid_c.cpp:264-289). mask_host may be NULL (= all envs). */
int mg_reset(void* handle, const uint32_t* seeds_host, const int32_t* map_index_host, const uint8_t* mask_host);

/* MettaGrid::step (mettagrid_c.cpp:572-678,740-752) for every env, including the auto-reset of
 * MettaGridPufferEnv.step (mettagrid/envs/mettagrid_puffer_env.py:158-164). Asynchronous on
 * `cuda_stream` (a cudaStream_t, may be NULL). */
int mg_step(void* handle, void* cuda_stream);

/* The same step for callers holding host (numpy) buffers, as MettaGridPufferEnv.step does:
 * copies actions in and observations/rewards/terminals/truncations out, then returns. */
int mg_step_host(void* handle, const int32_t* actions_host, uint8_t* obs_host, float* rewards_host,
                 uint8_t* terminals_host, uint8_t* truncations_host);

/* mg_step_host in two halves (mg_step_host = _begin, then _end): _begin enqueues the copies and the step and
 * returns at once, _end waits until the buffers given to _begin are filled; the buffers (and actions_host)
 * must stay untouched in between. For callers that split their envs over two handles and keep both in
 * flight, the way the reference's vectorised env overlaps its workers (pufferlib's send/recv with
 * batch_size = num_envs / 2, as metta/rl's async_factor 2 sets it up): one half's observations cross
 * the host link while the other half steps. One _begin per handle at a time. */
int mg_step_host_begin(void* handle, const int32_t* actions_host, uint8_t* obs_host, float* rewards_host,
                       uint8_t* terminals_host, uint8_t* truncations_host);
int mg_step_host_end(void* handle);

/* How mg_step_host fills a page-locked obs_host (no counterpart in the reference, whose
 * _compute_observations rewrites every byte of the caller's buffer each step, mettagrid_c.cpp:557-563,
 * 580). mode 0 (default): the whole buffer crosses the host link every call. mode 1: after one full
 * copy, a kernel writes only what can differ from what the buffer holds: of each row the leading
 * tokens (the rest of a row is 0xFF padding in both), and of those only the 64-byte pieces that differ
 * from a device-side copy of the buffer's contents (allocated by this call: one more observation
 * buffer in HBM). The buffer ends up byte-identical to mode 0 as long as nothing but this library
 * writes to it between calls. Another buffer address, or calling this again, starts over with a full
 * copy. Pageable buffers always take the full, staged copy. */
int mg_set_host_export(void* handle, int32_t mode);

/* Bytes mg_step_host has moved host->device and device->host through this handle so far. */
int mg_transfer_bytes(void* handle, uint64_t* h2d_bytes, uint64_t* d2h_bytes);

/* One synchronous step with CUDA events around each kernel (no counterpart in the reference; the
 * reference times its step with a host Stopwatch, mettagrid/simulator/simulator.py:138-160).
 * out_ms_host[0] = action kernel, out_ms_host[1] = observation kernel, in milliseconds. */
int mg_profile_step(void* handle, void* cuda_stream, float* out_ms_host);

/* MettaGrid::_compute_observations (mettagrid_c.cpp:557-563) alone. */
int mg_compute_observations(void* handle, void* cuda_stream);

/* What the policy's first op computes from the observations, handed over by the observation pass instead
 * (replaces the per-batch `coords == 255 -> argmax -> any -> max` of ObsShimTokens.forward,
 * agent/src/metta/agent/components/obs_shim.py:181-196). All three are optional caller-owned DEVICE
 * buffers (NULL = not produced), filled by every observation pass from then on:
 *   row_lengths i32[N*A]   tokens in the row = index of the first 0xFF location byte (T if the row is full)
 *   max_flip    i32[1]     max of row_lengths over all rows (after a masked mg_reset: an upper bound)
 *   obs_mask    u8[N*A][T] 1 where the slot is padding: the shim's `obs_mask`; slice [:, :max_flip]. */
int mg_set_obs_aux(void* handle, void* row_lengths_dev, void* max_flip_dev, void* obs_mask_dev);

/* auto_reset: 1 = MettaGridPufferEnv semantics (default), 0 = bare MettaGrid. map_stride: which pool
 * map an env moves to on auto-reset ((index + stride) mod pool size). */
int mg_set_options(void* handle, int32_t auto_reset, int32_t map_stride);

/* EarlyResetHandler (mettagrid/envs/early_reset_handler.py:12-20): truncate env i's running episode
 * once current_step >= steps[i] (0 = never). */
int mg_set_early_reset(void* handle, const uint32_t* steps_host);

/* State access for the conformance harness and for infos: the counterpart of grid_objects(),
 * get_episode_stats(), get_episode_rewards(), action_success() (mettagrid_c.cpp:754-957).
 * `what` selects a raw array (record layouts: metta_b200/csrc/mg_state.h):
 * 0 env scalars, 1 grid, 2 agents, 3 assemblers, 4 chests, 5 agent stats, 6 game stats,
 * 7 last executed action, 8 action success, 9 finished-episode accumulators, 10 rng state,
 * 11 observations, 12 rewards, 13 terminals, 14 truncations, 15 actions, 17 out-of-range action ids per agent
 * (i32 [N][A][MG_INVALID_WORDS], the record behind the action.invalid_index.<id> stats), 18 token records (what
 * every object currently shows to observers, i.e. its obs_features(): byte 0 = n, then n u16 feature | value << 8
 * from byte 8; layout from mg_token_record_layout). */
int64_t mg_sizeof(void* handle, int32_t what);
int mg_read(void* handle, int32_t what, void* out_host, int64_t nbytes);
int mg_write(void* handle, int32_t what, const void* in_host, int64_t nbytes);
/* out5: bytes per token record, records per env, first record of an env's agents, assemblers, chests (records
 * before that: one per map cell code). With mg_read what=18 this gives grid_objects() its feature entries
 * (mettagrid_c.cpp:800-806). */
int mg_token_record_layout(void* handle, int32_t* out5_host);

/* MettaGrid::set_inventory (mettagrid_c.cpp:959-965): items_amounts = n item ids followed by n amounts. */
int mg_set_inventory(void* handle, int32_t env, int32_t agent, const int32_t* items_amounts_host, int32_t n);

/* No counterpart in the reference (its core runs on the host and allocates freely).
 * Writes the observations of the CURRENT state of every env into obs_out_dev (u8[N*A][T][3], device
 * memory of the caller) without allocating anything: no cudaMalloc/cudaFree, no allocating library
 * call. For callers that must not allocate device memory while a rollout is running, e.g. a monitor
 * or replay thread peeking at a training run whose framework has reserved the whole GPU, or a
 * latency-bound thread that cannot take cudaMalloc's implicit synchronisation. Unlike
 * mg_compute_observations it leaves the handle's bound buffers and token statistics untouched: the
 * kernel's per-agent token-count scratch lives in a static 1 MiB device array instead.
 * Limits and rules: at most 262144 agents (N*A) per call, larger batches are rejected with an
 * error; every call on a GPU shares that array, which keeps the last call's contents (nothing in it
 * needs clearing), so calls that overlap on the GPU race and must be serialised by the caller; the
 * first call on each device looks up the array's device address (cudaGetSymbolAddress) and should
 * be made before any stream capture begins. Asynchronous on `cuda_stream`.
 * Ordering is the caller's job: the peek reads grid, agent and token records, so it has to run after
 * the preceding mg_step has finished and before the next one starts (same stream, or an event between
 * the two); on an unordered stream it would see a half-stepped state. Host calls on one handle must not
 * overlap either (the handle is not thread-safe); the peek itself writes nothing into the handle. */
int mg_peek_observations_noalloc(void* handle, void* obs_out_dev, void* cuda_stream);

int64_t mg_kernel_launches(void* handle); /* kernels launched through this handle so far */
int mg_sync(void* handle);

#ifdef __cplusplus
}
#endif
#endif /* METTAGRID_B200_H_ */
