"""GpuMettaGridVecEnv on a GPU: the pufferlib / TrainingEnvironment contract, episode-end infos
against the oracle's stats, host mode against device mode, and a rollout loop whose policy reads
the observation tensor the kernels write."""

import numpy as np
import pytest

from metta_b200.lowering import lower_game_config
from metta_b200.maps import names_to_codes
from metta_b200.stats import agent_stats_dict, game_stats_dict
from harness import random_actions
from scenarios import SCENARIOS

pytestmark = pytest.mark.gpu


def _make(name, n, **kw):
    from metta_b200.vecenv import GpuMettaGridVecEnv

    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    return cfg, low, codes, GpuMettaGridVecEnv(cfg, n, map_codes=codes[None], seed=11, **kw)


def test_contract_and_episode_infos_match_oracle():
    import torch

    from oracle.oracle import OracleEnv

    n = 6
    cfg, low, codes, env = _make("tiny_truncating", n)
    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    assert env.num_agents == n * A and env.single_observation_space.shape == (T, 3)
    assert env.single_action_space.n == len(low.action_names)
    obs, info = env.reset(seed=11)
    assert obs.shape == (n * A, T, 3) and obs.dtype == torch.uint8 and obs.is_cuda and info == {}
    orcs = [OracleEnv(low.words, codes, 11 + i) for i in range(n)]
    rng = np.random.RandomState(3)
    for step in range(1, 2 * cfg.game.max_steps + 1):
        a = random_actions(rng, cfg, low, n * A)
        env.send(torch.from_numpy(a).cuda())
        o, r, d, t, ta, infos, ids, mask = env.recv()
        if step <= cfg.game.max_steps:
            for i, orc in enumerate(orcs):
                orc.step(a[i * A : (i + 1) * A])
                assert np.array_equal(o[i * A : (i + 1) * A].cpu().numpy(), orc.observations)
        if step % cfg.game.max_steps == 0:
            assert bool(t.all()) and len(infos) == n
        else:
            assert not bool(t.any()) and infos == []
        if step == cfg.game.max_steps:
            groups = env.batch.read_agents()["group"]
            for i, (orc, inf) in enumerate(zip(orcs, infos)):
                assert inf["env_index"] == i and inf["attributes"]["steps"] == cfg.game.max_steps
                assert inf["attributes"]["seed"] == 11 + i
                assert inf["game"] == game_stats_dict(low, orc.game_stats(), codes)
                want: dict[str, float] = {}
                for ai in range(A):
                    for k, v in agent_stats_dict(low, int(groups[i, ai]), orc.agent_stats()[ai]).items():
                        want[k] = want.get(k, 0.0) + v
                want = {k: v / A for k, v in want.items()}
                assert inf["agent"].keys() == want.keys()
                for k in want:
                    assert inf["agent"][k] == pytest.approx(want[k], rel=1e-12), k
                assert np.array_equal(inf["episode_rewards"], orc.episode_rewards())
    got = env.get_observations()
    assert got[6] == slice(0, n * A) and got[8] == n * A
    env.close()


def test_host_mode_equals_device_mode():
    import torch

    n = 5
    cfg, low, codes, dev_env = _make("kitchen_sink", n)
    _, _, _, host_env = _make("kitchen_sink", n, host=True)
    A = cfg.game.num_agents
    dev_env.reset(seed=4)
    h_obs, _ = host_env.reset(seed=4)
    assert isinstance(h_obs, np.ndarray)
    assert np.array_equal(h_obs, dev_env.observations.cpu().numpy())
    rng = np.random.RandomState(8)
    for _ in range(30):
        a = random_actions(rng, cfg, low, n * A)
        o, r, d, t, _ = dev_env.step(torch.from_numpy(a).cuda())
        ho, hr, hd, ht, _ = host_env.step(a)
        assert np.array_equal(ho, o.cpu().numpy()) and np.array_equal(hr, r.cpu().numpy())
        assert np.array_equal(hd, d.cpu().numpy()) and np.array_equal(ht, t.cpu().numpy())
    dev_env.close()
    host_env.close()


def test_rollout_loop_with_a_torch_policy_reads_the_kernel_output_in_place():
    """configs[4] in small: env step and policy forward alternate on one stream with no host copy;
    the policy's input is the tensor the observation kernel writes."""
    import torch

    from metta_b200.builders import make_arena
    from metta_b200.maps import arena_maps
    from metta_b200.vecenv import GpuMettaGridVecEnv

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    env = GpuMettaGridVecEnv(cfg, 64, map_codes=arena_maps(low, 8, seed=5), seed=1)
    env.collect_infos = False
    obs, _ = env.reset()
    ptr = obs.data_ptr()
    torch.manual_seed(0)
    emb = torch.nn.Embedding(256, 16).cuda()
    head = torch.nn.Linear(16, env.single_action_space.n).cuda()
    total = torch.zeros((), device="cuda")
    with torch.no_grad():
        for _ in range(20):
            feats = emb(obs[:, :, 1].long()) * (obs[:, :, 0:1] != 255)   # feature-id embedding, padding masked
            logits = head(feats.sum(1))
            actions = torch.distributions.Categorical(logits=logits).sample().to(torch.int32)
            obs, rew, term, trunc, _ = env.step(actions)
            assert obs.data_ptr() == ptr
            total += rew.sum()
    assert torch.isfinite(total).item()
    st = env.batch.read_env()
    assert (st["current_step"] == 20).all()
    env.close()


def test_properties_and_config_swap():
    import torch

    n = 3
    cfg, low, codes, env = _make("tiny_truncating", n)
    assert (env.obs_width, env.obs_height) == (cfg.game.obs.width, cfg.game.obs.height)
    assert env.max_steps == cfg.game.max_steps and env.map_height == codes.shape[0] and env.map_width == codes.shape[1]
    assert env.resource_names == list(low.resource_names) and "wall" in env.object_type_names
    env.reset(seed=2)
    A = cfg.game.num_agents
    env.step(torch.zeros(n * A, dtype=torch.int32, device="cuda"))
    assert env.action_success().shape == (n, A) and (env.current_step == 1).all()
    # CurriculumEnv.set_mg_config: another task from the next reset on
    cfg2, grid2 = SCENARIOS["kitchen_sink"]()
    env.set_mg_config(cfg2, maps=[grid2])
    obs, _ = env.reset(seed=5)
    assert env.agents_per_env == cfg2.game.num_agents and obs.shape == (n * cfg2.game.num_agents, cfg2.game.obs.num_tokens, 3)
    _, _, _, ref = _make("kitchen_sink", n)
    ref_obs, _ = ref.reset(seed=5)
    assert torch.equal(obs, ref_obs)
    env.close()
    ref.close()


def test_obs_aux_matches_the_shim_expression_at_full_size():
    """(f3) row_lengths, max_flip and obs_mask written by the observation kernel equal what
    ObsShimTokens.forward computes per batch (obs_shim.py:181-196), on 4096 x 24 arena rows."""
    import torch

    from metta_b200.builders import make_arena
    from metta_b200.maps import arena_mapgen_pool
    from metta_b200.vecenv import GpuMettaGridVecEnv

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 4096
    env = GpuMettaGridVecEnv(cfg, n, map_codes=arena_mapgen_pool(low, 64), seed=42, obs_mask=True)
    env.collect_infos = False
    obs, _ = env.reset()
    g = torch.Generator().manual_seed(0)
    M = obs.shape[1]
    for step in range(12):
        pad = obs[..., 0] == 255
        flip = pad.int().argmax(dim=1)
        want = torch.where(pad.any(dim=1), flip, torch.full_like(flip, M))
        assert torch.equal(env.row_lengths.long(), want.long()), f"step {step}: row_lengths"
        assert int(env.max_flip) == int(want.max()), f"step {step}: max_flip"
        assert torch.equal(env.obs_mask, pad), f"step {step}: obs_mask"
        a = torch.randint(0, 5, (n * 24,), generator=g, dtype=torch.int32).cuda()
        obs, *_ = env.step(a)
    env.close()


def test_host_mode_uses_page_locked_arrays_and_matches_device_mode():
    """GpuMettaGridVecEnv(host=True) hands out page-locked numpy arrays filled through the changed-rows
    export; they must equal the device tensors of a twin env stepped with the same actions."""
    import torch

    cfg, low, codes, dev_env = _make("kitchen_sink", 40)
    _, _, _, host_env = _make("kitchen_sink", 40, host=True)
    A = cfg.game.num_agents
    o_d, _ = dev_env.reset(seed=5)
    o_h, _ = host_env.reset(seed=5)
    assert isinstance(o_h, np.ndarray) and torch.from_numpy(o_h).is_pinned()
    assert np.array_equal(o_h, o_d.cpu().numpy())
    rng = np.random.RandomState(9)
    for step in range(40):
        a = random_actions(rng, cfg, low, 40 * A)
        o_d, r_d, d_d, t_d, _ = dev_env.step(a)
        o_h, r_h, d_h, t_h, _ = host_env.step(a)
        assert np.array_equal(o_h, o_d.cpu().numpy()), f"step {step}: observations"
        assert np.array_equal(r_h, r_d.cpu().numpy()) and np.array_equal(d_h, d_d.cpu().numpy()) and np.array_equal(t_h, t_d.cpu().numpy())
    dev_env.close()
    host_env.close()


def test_grid_objects_dicts_match_the_compiled_reference():
    """GpuMettaGridVecEnv.grid_objects(): the reference's dicts (mettagrid_c.cpp:754-868) rebuilt from the
    device records, against the compiled reference stepping the same env."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("compiled reference (oracle/_ref) not present")
    from oracle.ref_bridge import RefEnv
    from test_kernels_emulated import _same_grid_objects

    cfg, low, codes, env = _make("kitchen_sink", 3)
    A = cfg.game.num_agents
    env.reset(seed=11)
    from metta_b200.maps import codes_to_names

    refs = [RefEnv(cfg.game, codes_to_names(low, codes), 11 + i) for i in range(3)]
    rng = np.random.RandomState(2)
    for s in range(50):
        a = random_actions(rng, cfg, low, 3 * A)
        env.step(a)
        for i, r in enumerate(refs):
            r.step(a[i * A : (i + 1) * A])
        if s % 10 == 9:
            for i, r in enumerate(refs):
                _same_grid_objects(r.grid_objects(), env.grid_objects(i))
    env.close()


def test_replay_matches_the_reference_writer():
    """metta_b200.replay on the GPU path against a replay the reference's own writer recorded (see harness)."""
    from harness import check_replay_against_reference

    check_replay_against_reference(cuda=True)


def test_replay_log_writer_writes_one_file_per_episode(tmp_path):
    """ReplayLogWriter beside a vec env: env 2 of 4 recorded over three episodes (auto-reset in between), one
    zlib-compressed JSON file per episode in the reference's layout."""
    import json
    import zlib

    import torch

    from metta_b200.replay import ReplayLogWriter

    n = 4
    cfg, low, codes, env = _make("tiny_truncating", n)
    A = cfg.game.num_agents
    env.reset(seed=11)
    writer = ReplayLogWriter(env, str(tmp_path), env_index=2)
    rng = np.random.RandomState(8)
    for step in range(3 * cfg.game.max_steps):
        a = torch.from_numpy(random_actions(rng, cfg, low, n * A)).cuda()
        env.step(a)
        writer.on_step(a)
    assert len(writer.paths) == 3
    for path in writer.paths:
        with open(path, "rb") as f:
            rep = json.loads(zlib.decompress(f.read()))
        assert rep["version"] == 3 and rep["max_steps"] == cfg.game.max_steps and rep["num_agents"] == A
        assert rep["action_names"] == low.action_names and rep["mg_config"]["game"]["num_agents"] == A
        agents = [o for o in rep["objects"] if o.get("is_agent")]
        assert len(agents) == A and all("location" in o and "total_reward" in o for o in agents)
    env.close()


def test_pair_of_halves_matches_one_env_object():
    """GpuMettaGridVecEnvPair (pufferlib's asynchronous contract with batch_size = num_envs / 2, the reference
    trainer's async_factor 2): the halves handed out in turn carry, step for step, what one host-mode env object
    over all envs returns for those envs -- observations, rewards, flags, agent ids, and the infos of finished
    episodes with their global env indices."""
    from metta_b200.vecenv import GpuMettaGridVecEnv, GpuMettaGridVecEnvPair

    cfg, grid = SCENARIOS["tiny_truncating"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    n, A = 6, cfg.game.num_agents
    one = GpuMettaGridVecEnv(cfg, n, map_codes=codes[None], seed=21, host=True)
    pair = GpuMettaGridVecEnvPair(cfg, n, map_codes=codes[None], seed=21)
    assert pair.batch_size == n * A // 2
    one.async_reset(21)
    pair.async_reset(21)
    rng = np.random.RandomState(5)
    episodes = 0
    for step in range(2 * cfg.game.max_steps + 3):
        o, r, d, t, _, infos, ids, _ = one.recv()
        acts = []
        for h in range(2):
            ph = pair.recv()
            sl = slice(h * n * A // 2, (h + 1) * n * A // 2)
            assert np.array_equal(ph[0], o[sl]), f"step {step} half {h}: observations"
            assert np.array_equal(ph[1], r[sl]) and np.array_equal(ph[2], d[sl]) and np.array_equal(ph[3], t[sl])
            assert np.array_equal(ph[6], ids[sl])
            want = [i for i in infos if (i["env_index"] >= n // 2) == bool(h)]
            assert [i["env_index"] for i in ph[5]] == [i["env_index"] for i in want]
            for got, w in zip(ph[5], want):
                assert got["game"] == w["game"] and got["agent"] == w["agent"] and got["attributes"]["seed"] == w["attributes"]["seed"]
            episodes += len(ph[5])
            a = random_actions(rng, cfg, low, n * A // 2)
            pair.send(a)
            acts.append(a)
        one.send(np.concatenate(acts))
    assert episodes == 2 * n
    one.close()
    pair.close()
