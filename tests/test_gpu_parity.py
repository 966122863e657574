"""GPU parity: the compiled sm_100a kernels, called through the C-ABI, against the oracle, the
reference-generated fixtures and (where the compiled reference travelled) the reference itself.
Bit-exact on every observation byte, reward, flag, inventory and stat."""

import numpy as np
import pytest
from golden_util import NAMES, load
from harness import HostBuffers, Lockstep, random_actions
from scenarios import SCENARIOS

from metta_b200.builders import make_arena
from metta_b200.batch import Batch, order_list
from metta_b200.layout import L
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_maps, build_pool, names_to_codes
from oracle import ref_loader

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(product_lib):
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return product_lib


@pytest.mark.parametrize("name", NAMES)
def test_cuda_reproduces_reference_fixture(name, lib):
    cfg, grid, d = load(name)
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    pool = build_pool(low, names_to_codes(low, grid))
    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, pool, 1)
    b.set_options(auto_reset=bool(d["restarts"]))
    buf = HostBuffers(A, T, cuda=True)
    b.set_buffers(*buf.ptrs)
    b.reset(np.array([int(d["seed"])], np.uint32), np.array([0], np.int32))
    obs, rew, term, trunc = buf.fetch()
    assert np.array_equal(obs, d["obs"][0]), "initial observation"
    for s, a in enumerate(d["actions"]):
        buf.set_actions(a)
        b.step()
        obs, rew, term, trunc = buf.fetch()
        assert np.array_equal(obs, d["obs"][s + 1]), f"{name}: observation at step {s}"
        assert np.array_equal(rew, d["rewards"][s]), f"{name}: rewards at step {s}"
        assert np.array_equal(term.astype(bool), d["terminals"][s]) and np.array_equal(trunc.astype(bool), d["truncations"][s])
        assert np.array_equal(b.read_action_success()[0], d["success"][s]), f"{name}: success at step {s}"
        assert np.array_equal(b.read_episode_rewards()[0], d["episode_rewards"][s])
    ag = b.read_agents()[0]
    assert np.array_equal(np.stack([ag["r"], ag["c"]], 1), d["final_pos"])
    for k in range(A):
        items = [int(x) for x in d["final_inv_items"][k] if x >= 0]
        assert order_list(ag["order"][k]) == items
        assert [int(ag["amt"][k][i]) for i in items] == [int(x) for x in d["final_inv_amounts"][k][: len(items)]]
    b.close()


@pytest.mark.parametrize("name,steps", [("kitchen_sink", 400), ("tiny_truncating", 40), ("clipped_start", 200), ("collective_rewards", 200)])
def test_scenarios_vs_oracle(name, steps, lib):
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    n = 33  # not a multiple of the warps per CTA: exercises the ragged last CTA
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=n, seeds=np.arange(n) + 100, cuda=True)
    ls.compare_outputs()
    rng = np.random.RandomState(5)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=3 if name == "collective_rewards" else 31, step=s))
        if name != "tiny_truncating" and s % 100 == 99:
            ls.compare_state()
    ls.close()


def test_arena_vs_oracle_with_state(lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 64
    ls = Lockstep(cfg, arena_maps(low, 16, seed=1), num_envs=n, seeds=np.arange(n) + 42, cuda=True)
    ls.compare_outputs()
    rng = np.random.RandomState(0)
    for s in range(300):
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.compare_state()
    ls.close()


def test_arena_auto_reset_and_map_rotation(lib):
    """Episodes end by truncation every 20 steps; envs restart from their pool map with the same seed."""
    cfg = make_arena(24, max_steps=20)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 62, 62)
    n = 8
    ls = Lockstep(cfg, arena_maps(low, 4, seed=2), num_envs=n, seeds=np.arange(n) + 7, cuda=True)
    rng = np.random.RandomState(3)
    for s in range(65):
        ls.step(rng.randint(0, 5, size=n * 24))
    env = ls.batch.read_env()
    assert np.all(env["episodes"] == 3) and np.all(env["current_step"] == 5)
    acc = ls.batch.read_episode_accum()
    assert acc[0] == 3 * n
    ls.close()


def test_rebinding_buffers_recaptures_the_step_graph(lib, monkeypatch):
    """set_buffers with other tensors in the middle of a run (the captured step graph holds the old
    addresses and has to be rebuilt), several slices."""
    monkeypatch.setenv("MG_CHUNKS", "4")
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 20
    ls = Lockstep(cfg, arena_maps(low, 3, seed=8), num_envs=n, seeds=np.arange(n) + 5, cuda=True)
    rng = np.random.RandomState(13)
    for s in range(30):
        if s in (10, 20):
            ls.buf = HostBuffers(ls.N * ls.A, ls.T, cuda=True)
            ls.batch.set_buffers(*ls.buf.ptrs)
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.close()


def test_queue_mode_matches_oracle(lib, monkeypatch):
    """MG_STEP_MODE=queue: step kernel with one env per CTA handing finished envs to observation CTAs
    that run beside it. Same results as the default sliced launch."""
    monkeypatch.setenv("MG_STEP_MODE", "queue")
    cfg = make_arena(24, max_steps=25)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 62, 62)
    n = 40
    ls = Lockstep(cfg, arena_maps(low, 4, seed=6), num_envs=n, seeds=np.arange(n) + 70, cuda=True)
    rng = np.random.RandomState(9)
    for s in range(60):
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.compare_state()
    ls.batch.sync()
    ls.close()


@pytest.mark.parametrize("n", [40, 1500])
def test_flag_mode_matches_oracle(n, lib, monkeypatch):
    """MG_STEP_MODE=flags: the observation kernel runs beside the step kernel and every CTA waits for its
    env's mark. Same results as the default sliced launch, across episode ends (auto-reset), also with more
    envs than fit beside each other on the GPU's SMs at the observation kernel's CTA size."""
    monkeypatch.setenv("MG_STEP_MODE", "flags")
    cfg = make_arena(24, max_steps=25)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 62, 62)
    ls = Lockstep(cfg, arena_maps(low, 4, seed=6), num_envs=n, seeds=np.arange(n) + 70, cuda=True)
    rng = np.random.RandomState(9)
    for s in range(60 if n < 100 else 8):
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.compare_state()
    ls.batch.sync()
    ls.close()


@pytest.mark.skipif(not ref_loader.available(), reason="compiled reference (oracle/_ref) not present")
def test_arena_vs_compiled_reference(lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 6
    ls = Lockstep(cfg, arena_maps(low, n, seed=11), num_envs=n, seeds=np.arange(n) + 1, cuda=True, checker="reference")
    ls.compare_outputs()
    rng = np.random.RandomState(4)
    for s in range(200):
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.close()


def test_full_size_properties(lib):
    """BASELINE config 2 (4096 envs x 24 agents): 1000 steps in lockstep with the oracle on a sample of
    envs (outputs compared every 50 steps and at the end), plus size-independent checks over all envs."""
    import torch

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n, A, T = 4096, 24, 200
    codes = arena_maps(low, 256, seed=0)
    pool = build_pool(low, codes)
    b = Batch(low, pool, n)
    buf = HostBuffers(n * A, T, cuda=True)
    b.set_buffers(*buf.ptrs)
    seeds = (np.arange(n) + 42).astype(np.uint32)
    maps = (np.arange(n) % 256).astype(np.int32)
    b.reset(seeds, maps)
    from oracle.oracle import OracleEnv

    sample = [0, 1, 255, 256, 2047, 4095]
    orcs = {i: OracleEnv(low.words, codes[maps[i]], int(seeds[i])) for i in sample}
    g = torch.Generator().manual_seed(0)
    for s in range(1000):
        a = torch.randint(0, 5, (n * A,), generator=g, dtype=torch.int32)
        buf.t_act.copy_(a)
        b.step()
        an = a.numpy().reshape(n, A)
        for i, o in orcs.items():
            o.step(an[i])
        if s % 50 == 49:
            obs, rew, term, trunc = buf.fetch()
            for i, o in orcs.items():
                assert np.array_equal(obs[i * A : (i + 1) * A], o.observations), f"step {s} env {i}"
                assert np.array_equal(rew[i * A : (i + 1) * A], o.rewards), f"step {s} env {i}"
    obs, rew, term, trunc = buf.fetch()
    # properties over all 98304 agents
    grid = b.read_grid()
    ag = b.read_agents()
    kinds = grid >> 14
    assert np.all((kinds == 1).sum(axis=(1, 2)) == A), "every env still has exactly A agents on its grid"
    e = np.repeat(np.arange(n), A)
    k = np.tile(np.arange(A), n)
    assert np.array_equal(grid[e, ag["r"].reshape(-1), ag["c"].reshape(-1)], (0x4000 | k).astype(np.uint16)), "cell<->agent consistency"
    o3 = obs.reshape(n * A, T, 3)
    empty = (o3 == 255).all(axis=2)
    assert np.all(np.diff(empty.astype(np.int8), axis=1) >= 0), "written tokens form a prefix, the rest is 0xFF"
    assert np.all(o3[:, 0, 0] == 0x55) and np.all(o3[:, 0, 1] == low.feature_ids["episode_completion_pct"])
    # same (map, seed, actions) => same result: env i and env i+256 share a map; re-run env 0's twin
    gs = b.read_game_stats()
    written = (~empty).sum(axis=1).reshape(n, A).sum(axis=1)
    assert np.all(gs[:, L.MG_GS_TOKENS_WRITTEN] > 0) and written.min() > 0
    b.close()


@pytest.mark.parametrize("pinned,changed_only", [(False, False), (True, False), (True, True)])
def test_host_buffer_step_matches_oracle(pinned, changed_only, lib):
    """mg_step_host (numpy in, numpy out) with pageable and with page-locked buffers, and with the
    observation export that only writes the part of each row that changed."""
    import torch

    from oracle.oracle import OracleEnv

    cfg, grid = SCENARIOS["kitchen_sink"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    n, A, T = 5, cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, build_pool(low, codes[None]), n)
    b.use_internal_buffers()
    b.set_host_export(changed_only)
    seeds = np.arange(n, dtype=np.uint32) + 9
    b.reset(seeds, np.zeros(n, np.int32))
    mk = (lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()) if pinned else (
        lambda shape, dt: np.empty(shape, dtype=torch.empty(0, dtype=dt).numpy().dtype))
    act, obs = mk((n * A,), torch.int32), mk((n * A, T, 3), torch.uint8)
    rew, term, trunc = mk((n * A,), torch.float32), mk((n * A,), torch.uint8), mk((n * A,), torch.uint8)
    orcs = [OracleEnv(low.words, codes, int(s)) for s in seeds]
    rng = np.random.RandomState(12)
    for s in range(40):
        act[:] = random_actions(rng, cfg, low, n * A)
        b.step_host(act, obs, rew, term, trunc)
        for i, o in enumerate(orcs):
            o.step(act[i * A : (i + 1) * A])
            assert np.array_equal(obs[i * A : (i + 1) * A], o.observations), f"step {s} env {i}"
            assert np.array_equal(rew[i * A : (i + 1) * A], o.rewards)
    h2d, d2h = b.transfer_bytes()
    full = 40 * n * A * (T * 3 + 6)
    assert h2d == 40 * n * A * 4
    # (the export widens each row's prefix to whole 32-byte sectors, so with rows this short it need not be smaller)
    assert (0 < d2h != full) if changed_only else (d2h == full), (d2h, full)
    b.close()


@pytest.mark.parametrize("pinned", [False, True])
def test_two_handles_in_flight_match_oracle(pinned, lib):
    """mg_step_host_begin / mg_step_host_end: two handles (two halves of a caller's envs) stepped as a
    software pipeline -- one half's results are read while the other half is in flight -- give what the
    oracle gives for each half; a second _begin without _end and an _end without _begin are refused."""
    import torch

    from metta_b200.capi import MgError
    from oracle.oracle import OracleEnv

    cfg, grid = SCENARIOS["kitchen_sink"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    n, A, T = 6, cfg.game.num_agents, cfg.game.obs.num_tokens
    mk = (lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()) if pinned else (
        lambda shape, dt: np.empty(shape, dtype=torch.empty(0, dtype=dt).numpy().dtype))
    halves = []
    for h in range(2):
        b = Batch(low, build_pool(low, codes[None]), n)
        b.use_internal_buffers()
        b.set_host_export(pinned)
        seeds = np.arange(n, dtype=np.uint32) + 100 * (h + 1)
        b.reset(seeds, np.zeros(n, np.int32))
        bufs = (mk((n * A, T, 3), torch.uint8), mk((n * A,), torch.float32), mk((n * A,), torch.uint8), mk((n * A,), torch.uint8))
        halves.append((b, bufs, mk((n * A,), torch.int32), [OracleEnv(low.words, codes, int(s)) for s in seeds]))
    with pytest.raises(MgError):
        halves[0][0].step_host_end()
    rng = np.random.RandomState(3)
    sent = [None, None]
    for h, (b, bufs, act, _) in enumerate(halves):
        act[:] = random_actions(rng, cfg, low, n * A)
        sent[h] = act.copy()
        b.step_host_begin(act, *bufs)
    with pytest.raises(MgError):
        halves[0][0].step_host_begin(halves[0][2], *halves[0][1])
    for s in range(30):
        for h, (b, bufs, act, orcs) in enumerate(halves):
            b.step_host_end()
            for i, o in enumerate(orcs):
                o.step(sent[h][i * A : (i + 1) * A])
                assert np.array_equal(bufs[0][i * A : (i + 1) * A], o.observations), f"step {s} half {h} env {i}"
                assert np.array_equal(bufs[1][i * A : (i + 1) * A], o.rewards)
            act[:] = random_actions(rng, cfg, low, n * A)
            sent[h] = act.copy()
            b.step_host_begin(act, *bufs)
    for b, _, _, _ in halves:
        b.step_host_end()
        b.close()


@pytest.mark.parametrize("chunks", ["1", "3"])
def test_changed_rows_export_follows_resets(chunks, lib, monkeypatch):
    """Export mode 1 across episode ends (token counts drop back at a reset), a second buffer and a
    mode switch: the page-locked host buffer always equals the device observations. With MG_CHUNKS=3
    the export kernels run inside the captured step graph, one per slice of envs."""
    import torch

    monkeypatch.setenv("MG_CHUNKS", chunks)

    cfg, grid = SCENARIOS["tiny_truncating"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    n, A, T = 67, cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, build_pool(low, codes[None]), n)
    b.use_internal_buffers()
    b.set_host_export(True)
    b.reset(np.arange(n, dtype=np.uint32) + 3, np.zeros(n, np.int32))
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()  # noqa: E731
    act = pin((n * A,), torch.int32)
    obs_a, obs_b = pin((n * A, T, 3), torch.uint8), pin((n * A, T, 3), torch.uint8)
    obs_a[...] = 7  # garbage: the first call on a buffer must not rely on its contents
    obs_b[...] = 9
    rew, term, trunc = pin((n * A,), torch.float32), pin((n * A,), torch.uint8), pin((n * A,), torch.uint8)
    rng = np.random.RandomState(5)
    episodes_seen = 0
    for s in range(3 * cfg.game.max_steps + 5):
        act[:] = random_actions(rng, cfg, low, n * A)
        buf = obs_b if 20 <= s < 30 else obs_a
        if s == 40:
            b.set_host_export(False)
        if s == 45:
            b.set_host_export(True)
        b.step_host(act, buf, rew, term, trunc)
        dev = b.read_obs()
        assert np.array_equal(buf, dev.reshape(buf.shape)), f"step {s}"
        episodes_seen += int(trunc.any() or term.any())
    assert episodes_seen >= 2
    b.close()


@pytest.mark.parametrize("name,n,steps", [("dense_conflict", 64, 2000), ("crowd_48", 32, 600)])
def test_crowded_rooms_vs_oracle(name, n, steps, lib):
    """Where the reference's shuffled processing order decides most outcomes: 28 agents in an 8x7 room
    (attack, transfer, swaps with frozen agents, shared assembler neighbourhoods, chests in reach of an
    assembler, loot that removes a limit modifier), and 48 agents (the 64-agent instantiation of the
    step kernel). The oracle is pinned against the compiled reference on the same scenarios
    (tests/test_oracle_vs_reference.py::test_crowded_rooms_many_seeds)."""
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=n, seeds=np.arange(n) + 300, cuda=True)
    ls.compare_outputs()
    rng = np.random.RandomState(17)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=41, step=s))
        if s % 50 == 49:
            ls.compare_state()
    ordered = ls.batch.read_env()["spare"][:, 0].sum() / float(n * ls.A * steps)
    st = ls.batch.read_agent_stats().sum(axis=(0, 1))
    assert st[L.MG_AS_SWAP] > 0 and st[L.MG_AS_TRANSFER_COUNT] > 0 and st[L.MG_AS_ATTACK_HIT : L.MG_AS_ATTACK_HIT + 8].sum() > 0
    assert ordered > 0.6, f"only {ordered:.2f} of the agents went through the ordered pass"
    print(f"{name}: {ordered:.3f} of all agent-steps went through the ordered pass")
    ls.close()


def test_shuffle_with_a_rejected_word(lib):
    """Lemire's method in std::shuffle rejects about one word in 1e7 (bits/uniform_int_dist.h:257-330); the
    kernel then leaves its draw-per-lane path for the serial loop. Forced here by zeroing the state word of
    the second draw (a zero word tempers to zero, and 0 * 12 mod 2^32 < 2^32 mod 12) in both generators."""
    cfg, grid = SCENARIOS["dense_conflict"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    n = 9
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=n, seeds=np.arange(n) + 900, cuda=True)
    rng = np.random.RandomState(23)
    for s in range(40):
        if s % 4 == 1:
            state = ls.batch.read_rng()
            idx = ls.batch.read_env()["rng_idx"].astype(np.int64)
            for i, o in enumerate(ls.checkers):
                mine = o.get_rng()
                assert int(mine[624]) == int(idx[i]) and np.array_equal(mine[:624], state[i]), "generators differ before the test"
                if idx[i] + 2 >= 624:
                    continue
                state[i, idx[i] + 1] = 0
                mine[idx[i] + 1] = 0
                o.set_rng(mine)
            before = idx.copy()
            ls.batch._write(10, state)
            ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
            after = ls.batch.read_env()["rng_idx"].astype(np.int64)
            for i, o in enumerate(ls.checkers):
                assert int(o.get_rng()[624]) == int(after[i])
                if before[i] + 2 < 624 and after[i] > before[i]:
                    assert after[i] - before[i] >= ls.A // 2 + 1, "the rejected word was not replaced by another draw"
        else:
            ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
    ls.compare_state()
    ls.close()


def test_reset_while_a_step_is_in_flight_and_partial_reset(lib):
    """mg_reset right behind an asynchronous mg_step on a non-blocking stream (it must wait for the step),
    with a mask: unmasked envs keep running their episode."""
    import torch

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 24
    ls = Lockstep(cfg, arena_maps(low, 4, seed=4), num_envs=n, seeds=np.arange(n) + 11, cuda=True)
    rng = np.random.RandomState(6)
    side = torch.cuda.Stream()
    for s in range(12):
        ls.step(rng.randint(0, 5, size=n * 24))
    a = rng.randint(0, 5, size=(n, 24)).astype(np.int32)
    ls.buf.set_actions(a.reshape(-1))
    torch.cuda.synchronize()
    ls.batch.step(side.cuda_stream)  # in flight on a side stream
    mask = (np.arange(n) % 3 == 0).astype(np.uint8)
    new_seeds = ls.seeds.copy()
    new_seeds[mask.astype(bool)] += 1000
    ls.batch.reset(new_seeds, ls.map_of_env.astype(np.int32), mask)
    ls.step_no += 1
    for i, ck in enumerate(ls.checkers):
        ck.step(a[i])
        if mask[i]:
            ls.seeds[i] = new_seeds[i]
            ls.checkers[i] = ls._make_checker(i)
    ls.compare_outputs()
    for s in range(10):
        ls.step(rng.randint(0, 5, size=n * 24))
    ls.compare_state()
    ls.close()


def test_handle_on_a_device_that_is_not_current(lib):
    """Every entry point makes its handle's device current itself: a batch on cuda:1 steps correctly while
    cuda:0 is the caller's current device (needs two GPUs)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    cfg, grid = SCENARIOS["kitchen_sink"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    from oracle.oracle import OracleEnv

    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    torch.cuda.set_device(0)
    b = Batch(low, build_pool(low, codes), 2, device=1)
    obs = torch.zeros((2 * A, T, 3), dtype=torch.uint8, device="cuda:1")
    term = torch.zeros(2 * A, dtype=torch.uint8, device="cuda:1")
    trunc = torch.zeros_like(term)
    rew = torch.zeros(2 * A, dtype=torch.float32, device="cuda:1")
    act = torch.zeros(2 * A, dtype=torch.int32, device="cuda:1")
    b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
    b.reset(np.array([5, 6], np.uint32), np.zeros(2, np.int32))
    orcs = [OracleEnv(low.words, codes, 5 + i) for i in range(2)]
    rng = np.random.RandomState(2)
    for s in range(20):
        a = random_actions(rng, cfg, low, 2 * A)
        act.copy_(torch.from_numpy(a))
        torch.cuda.synchronize(1)
        torch.cuda.set_device(0)
        b.step(0)
        b.sync()
        for i, o in enumerate(orcs):
            o.step(a[i * A : (i + 1) * A])
            assert np.array_equal(obs[i * A : (i + 1) * A].cpu().numpy(), o.observations), f"step {s} env {i}"
    b.close()


@pytest.mark.parametrize("name,n,steps", [("arena", 40, 60), ("kitchen_sink", 33, 200), ("crowd_48", 9, 120), ("dense_conflict", 70, 300)])
def test_fused_mode_matches_oracle(name, n, steps, lib, monkeypatch):
    """MG_STEP_MODE=fused: one persistent kernel whose warps step envs and then encode the observations of
    finished envs. Same results as the sliced launch, for the 32- and 64-agent instantiations and both
    window sizes, with episode ends inside the run."""
    monkeypatch.setenv("MG_STEP_MODE", "fused")
    if name == "arena":
        cfg = make_arena(24, max_steps=25)
        cfg.game.episode_truncates = True
        low = lower_game_config(cfg.game, 62, 62)
        codes = arena_maps(low, 4, seed=6)
    else:
        cfg, grid = SCENARIOS[name]()
        low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
        codes = names_to_codes(low, grid)
    ls = Lockstep(cfg, codes, num_envs=n, seeds=np.arange(n) + 70, cuda=True)
    ls.compare_outputs()
    rng = np.random.RandomState(9)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=37, step=s))
    ls.compare_state()
    ls.batch.sync()
    ls.close()


def test_token_stats_beyond_2_pow_24(lib):
    """tokens_free_space passes 2^24 after ~4700 arena steps; from there the reference's float additions
    round agent by agent (mettagrid_c.cpp:552-554) and the kernel replays them in that order."""
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    n = 3
    ls = Lockstep(cfg, arena_maps(low, 3, seed=12), num_envs=n, seeds=np.arange(n) + 61, cuda=True)
    rng = np.random.RandomState(14)
    for s in range(5600):
        ls.step(rng.randint(0, 5, size=n * 24), compare=(s % 400 == 0))
    ls.compare_outputs()
    ls.compare_state()
    assert ls.batch.read_game_stats()[:, L.MG_GS_TOKENS_FREE_SPACE].min() > 2.0**24
    ls.close()


def test_library_owned_buffers_are_rebound_by_every_all_null_call(lib):
    """mg_set_buffers(all NULL) -> caller buffers -> all NULL again: the third call must bind the library's own
    buffers again (it used to leave the handle pointing at the caller's, possibly freed, memory)."""
    import torch

    cfg, grid = SCENARIOS["tiny_truncating"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    n, A, T = 4, cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, build_pool(low, codes), n)
    b.use_internal_buffers()
    b.reset(np.arange(n, dtype=np.uint32) + 1, np.zeros(n, np.int32))
    own_first = b.read_obs().copy()
    buf = HostBuffers(n * A, T, cuda=True)
    b.set_buffers(*buf.ptrs)
    b.compute_observations()
    torch.cuda.synchronize()
    assert np.array_equal(buf.fetch()[0], own_first)
    ptr = buf.t_obs.data_ptr()
    b.use_internal_buffers()
    h = lambda shape, dt: np.zeros(shape, dt)  # noqa: E731
    obs, rew, term, trunc = h((n * A, T, 3), np.uint8), h(n * A, np.float32), h(n * A, np.uint8), h(n * A, np.uint8)
    buf.t_obs.fill_(7)  # the caller's buffer must not be touched any more
    b.step_host(np.zeros(n * A, np.int32), obs, rew, term, trunc)
    assert int(buf.t_obs.min()) == 7 and buf.t_obs.data_ptr() == ptr
    assert np.array_equal(obs, b.read_obs())
    b.close()
