"""The oracle (oracle/mg_oracle.c) against fixtures produced by the reference itself
(tests/golden/make_golden.py): every observation byte, reward, flag, action_success, episode
reward, and the final inventories in the reference's iteration order."""

import numpy as np
import pytest
from golden_util import NAMES, load

from metta_b200.lowering import lower_game_config
from metta_b200.maps import names_to_codes
from oracle.oracle import OracleEnv


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_reference_fixture(name, product_lib):
    cfg, grid, d = load(name)
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    seed = int(d["seed"])
    env = OracleEnv(low.words, codes, seed)
    assert np.array_equal(env.observations, d["obs"][0]), "initial observation"
    for s, a in enumerate(d["actions"]):
        if bool(d["restarts"]) and (env.terminals.all() or env.truncations.all()):
            env = OracleEnv(low.words, codes, seed)
        env.step(a)
        assert np.array_equal(env.observations, d["obs"][s + 1]), f"{name}: observation at step {s}"
        assert np.array_equal(env.rewards, d["rewards"][s]), f"{name}: rewards at step {s}"
        assert np.array_equal(env.terminals, d["terminals"][s]) and np.array_equal(env.truncations, d["truncations"][s])
        assert np.array_equal(env.action_success(), d["success"][s]), f"{name}: action_success at step {s}"
        assert np.array_equal(env.episode_rewards(), d["episode_rewards"][s])
    st = env.agent_state()
    assert np.array_equal(st[:, :2], d["final_pos"])
    for k in range(cfg.game.num_agents):
        items = [int(x) for x in d["final_inv_items"][k] if x >= 0]
        assert [int(x) for x in st[k, 6:22] if x >= 0] == items, "inventory iteration order"
        assert [int(st[k, 22 + i]) for i in items] == [int(x) for x in d["final_inv_amounts"][k][: len(items)]]
