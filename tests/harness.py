"""Lockstep conformance harness: one Batch (CUDA, or the emulated kernels) against per-env checkers."""

from __future__ import annotations

import numpy as np

from metta_b200.batch import Batch, order_list
from metta_b200.layout import L
from metta_b200.lowering import lower_game_config
from metta_b200.maps import build_pool, codes_to_names, names_to_codes
from oracle.oracle import OracleEnv


class HostBuffers:
    """Caller-owned buffers. With the emulator they are the 'device' buffers themselves; on a GPU
    they mirror torch CUDA tensors."""

    def __init__(self, n_agents: int, T: int, cuda: bool):
        self.cuda = cuda
        if cuda:
            import torch

            self.t_obs = torch.zeros((n_agents, T, 3), dtype=torch.uint8, device="cuda")
            self.t_term = torch.zeros(n_agents, dtype=torch.uint8, device="cuda")
            self.t_trunc = torch.zeros(n_agents, dtype=torch.uint8, device="cuda")
            self.t_rew = torch.zeros(n_agents, dtype=torch.float32, device="cuda")
            self.t_act = torch.zeros(n_agents, dtype=torch.int32, device="cuda")
            self.ptrs = [t.data_ptr() for t in (self.t_obs, self.t_term, self.t_trunc, self.t_rew, self.t_act)]
        else:
            self.obs = np.zeros((n_agents, T, 3), np.uint8)
            self.term = np.zeros(n_agents, np.uint8)
            self.trunc = np.zeros(n_agents, np.uint8)
            self.rew = np.zeros(n_agents, np.float32)
            self.act = np.zeros(n_agents, np.int32)
            self.ptrs = [a.ctypes.data for a in (self.obs, self.term, self.trunc, self.rew, self.act)]

    def set_actions(self, a: np.ndarray) -> None:
        if self.cuda:
            import torch

            self.t_act.copy_(torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)))
        else:
            self.act[:] = a

    def fetch(self):
        if self.cuda:
            import torch

            torch.cuda.synchronize()
            return (self.t_obs.cpu().numpy(), self.t_rew.cpu().numpy(), self.t_term.cpu().numpy(), self.t_trunc.cpu().numpy())
        return self.obs, self.rew, self.term, self.trunc


class Lockstep:
    def __init__(self, cfg, codes: np.ndarray, num_envs: int, seeds, clib=None, cuda: bool = False, checker="oracle",
                 map_of_env=None, auto_reset: bool = True):
        game = cfg.game
        codes = np.asarray(codes, dtype=np.uint8)
        if codes.ndim == 2:
            codes = codes[None]
        self.cfg, self.codes = cfg, codes
        self.lowered = lower_game_config(game, codes.shape[1], codes.shape[2])
        self.pool = build_pool(self.lowered, codes)
        self.N, self.A, self.T = num_envs, game.num_agents, game.obs.num_tokens
        self.batch = Batch(self.lowered, self.pool, num_envs, clib=clib)
        self.batch.set_options(auto_reset=auto_reset, map_stride=0)
        self.buf = HostBuffers(self.N * self.A, self.T, cuda)
        self.batch.set_buffers(*self.buf.ptrs)
        self.seeds = np.asarray(seeds, dtype=np.uint32)
        self.map_of_env = np.arange(num_envs) % len(codes) if map_of_env is None else np.asarray(map_of_env)
        self.batch.reset(self.seeds, self.map_of_env.astype(np.int32))
        self.checker_kind = checker
        self.auto_reset = auto_reset
        self.checkers = [self._make_checker(i) for i in range(num_envs)]
        self.step_no = 0

    def _make_checker(self, i: int):
        codes = self.codes[self.map_of_env[i]]
        if self.checker_kind == "oracle":
            return OracleEnv(self.lowered.words, codes, int(self.seeds[i]))
        from oracle.ref_bridge import RefEnv

        return RefEnv(self.cfg.game, codes_to_names(self.lowered, codes), int(self.seeds[i]))

    def compare_outputs(self) -> None:
        obs, rew, term, trunc = self.buf.fetch()
        A = self.A
        for i, ck in enumerate(self.checkers):
            sl = slice(i * A, (i + 1) * A)
            if not np.array_equal(obs[sl], ck.observations):
                d = np.argwhere(obs[sl] != ck.observations)
                a = int(d[0][0])
                raise AssertionError(
                    f"step {self.step_no} env {i} agent {a}: observation differs at token {int(d[0][1])}\n"
                    f"  sim     {obs[sl][a, :40].tolist()}\n  checker {ck.observations[a, :40].tolist()}"
                )
            assert np.array_equal(rew[sl], ck.rewards), f"step {self.step_no} env {i}: rewards {rew[sl]} vs {ck.rewards}"
            assert np.array_equal(term[sl].astype(bool), ck.terminals), f"step {self.step_no} env {i}: terminals"
            assert np.array_equal(trunc[sl].astype(bool), ck.truncations), f"step {self.step_no} env {i}: truncations"

    def step(self, actions: np.ndarray, compare: bool = True) -> None:
        actions = np.ascontiguousarray(actions, dtype=np.int32).reshape(self.N, self.A)
        self.buf.set_actions(actions.reshape(-1))
        self.batch.step()
        self.step_no += 1
        for i, ck in enumerate(self.checkers):
            if self.auto_reset and (ck.terminals.all() or ck.truncations.all()):
                # MettaGridPufferEnv.step (mettagrid_puffer_env.py:158-164): new sim, same seed, then step
                ck = self.checkers[i] = self._make_checker(i)
            ck.step(actions[i])
        if compare:
            self.compare_outputs()

    def compare_state(self) -> None:
        """Positions, vibes, frozen, inventories (with iteration order), stats, assembler/chest state."""
        assert self.checker_kind == "oracle"
        ag = self.batch.read_agents()
        st = self.batch.read_agent_stats()
        gs = self.batch.read_game_stats()
        asm = self.batch.read_assemblers()
        ch = self.batch.read_chests()
        succ = self.batch.read_action_success()
        inval = self.batch.read_invalid_ids()
        RS = L.MG_RES_STRIDE
        for i, o in enumerate(self.checkers):
            s = o.agent_state()
            assert np.array_equal(ag[i]["r"], s[:, 0]) and np.array_equal(ag[i]["c"], s[:, 1]), f"env {i}: positions"
            assert np.array_equal(ag[i]["vibe"], s[:, 2]) and np.array_equal(ag[i]["frozen"], s[:, 3]), f"env {i}: vibe/frozen"
            assert np.array_equal(ag[i]["steps_without_motion"], s[:, 4].astype(np.uint32))
            assert np.array_equal(ag[i]["amt"], s[:, 6 + RS : 6 + RS + L.MG_MAX_RESOURCES]), f"env {i}: inventories"
            for k in range(self.A):
                assert order_list(ag[i]["order"][k]) == [int(x) for x in s[k, 6 : 6 + RS] if x >= 0], f"env {i} agent {k}: order"
            assert np.array_equal(ag[i]["episode_reward"], o.episode_rewards()), f"env {i}: episode rewards"
            a_ref = o.agent_stats()
            assert np.array_equal(st[i], a_ref), f"env {i}: agent stats differ at {np.argwhere(st[i] != a_ref)[:4].tolist()}"
            g_ref = o.game_stats()
            assert np.array_equal(gs[i], g_ref), f"env {i}: game stats differ at {np.argwhere(gs[i] != g_ref)[:4].tolist()}"
            assert np.array_equal(succ[i], o.action_success()), f"env {i}: action_success"
            assert np.array_equal(inval[i], o.invalid_ids()), f"env {i}: record of out-of-range action ids"
            os_ = o.assembler_state()
            n = len(os_)
            if n:
                assert np.array_equal(asm[i]["r"][:n], os_[:, 0]) and np.array_equal(asm[i]["c"][:n], os_[:, 1])
                assert np.array_equal(asm[i]["flags"][:n] & 1, os_[:, 2]), f"env {i}: clipped flags"
                assert np.array_equal(asm[i]["cooldown_end"][:n], os_[:, 4].astype(np.uint32)), f"env {i}: cooldown_end"
                assert np.array_equal(asm[i]["uses_count"][:n], os_[:, 6].astype(np.uint32)), f"env {i}: uses_count"
                clipped = os_[:, 2] != 0
                assert np.array_equal(asm[i]["unclip_protocol"][:n][clipped], os_[:, 3][clipped]), f"env {i}: unclip protocol"
            oc = o.chest_state()
            for k in range(len(oc)):
                assert np.array_equal(ch[i]["amt"][k], oc[k, 3 + RS : 3 + RS + L.MG_MAX_RESOURCES]), f"env {i} chest {k}"
                assert order_list(ch[i]["order"][k]) == [int(x) for x in oc[k, 3 : 3 + RS] if x >= 0]

    def close(self):
        self.batch.close()


def random_actions(rng: np.random.RandomState, cfg, lowered, n: int, invalid_every: int = 0, step: int = 0) -> np.ndarray:
    """Random actions biased toward moves (so agents interact), optionally with out-of-range ids."""
    nact = len(lowered.action_names)
    nmove = len(cfg.game.actions.move.allowed_directions)
    a = rng.randint(0, nact, size=n).astype(np.int32)
    m = rng.rand(n) < 0.6
    a[m] = rng.randint(1, 1 + nmove, size=int(m.sum()))
    if invalid_every and step % invalid_every == 0:
        a[rng.randint(0, n)] = nact + rng.randint(0, 3)
        a[rng.randint(0, n)] = -1
    return a


def check_replay_against_reference(clib=None, cuda: bool = False) -> None:
    """tests/golden/arena_replay.*: an arena episode of 60 steps recorded by the reference's own EpisodeReplay
    (tests/golden/make_golden.py replay). The same episode stepped here (env 1 of 2, so that the env index is
    exercised) and recorded by metta_b200.replay must give the same replay: header, object records, every
    change list -- everything but the embedded config dump, which is each side's own schema object."""
    import json
    import zlib

    from golden_util import GOLDEN, load

    from metta_b200.maps import names_to_codes
    from metta_b200.replay import EpisodeReplay

    cfg, grid, data = load("arena_replay")
    with open(GOLDEN + "/arena_replay.replay.json.z", "rb") as f:
        want = json.loads(zlib.decompress(f.read()))
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    seed = int(data["seed"])
    ls = Lockstep(cfg, codes, num_envs=2, seeds=[seed + 5, seed], clib=clib, cuda=cuda)
    rep = EpisodeReplay(ls.batch, 1, mg_config=None)
    A = ls.A
    rng = np.random.RandomState(4)
    for s, a in enumerate(data["actions"]):
        ls.step(np.concatenate([rng.randint(0, 5, size=A), a]))
        _, rew, _, _ = ls.buf.fetch()
        rep.log_step(s + 1, a, rew[A : 2 * A], ls.batch.read_action_success()[1])
    got = json.loads(json.dumps(rep.data()))  # tuples -> lists, as the file would hold them
    ls.close()
    for key in want:
        if key in ("mg_config", "objects"):
            continue
        assert got[key] == want[key], f"replay header {key}: {got[key]} vs {want[key]}"
    assert len(got["objects"]) == len(want["objects"])
    for i, (g, w) in enumerate(zip(got["objects"], want["objects"])):
        assert g.keys() == w.keys(), f"object {i}: keys {sorted(g.keys() ^ w.keys())}"
        for k in w:
            assert g[k] == w[k], f"object {i} ({w.get('type_name')}) {k}: {g[k]} vs {w[k]}"
