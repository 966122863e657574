"""Runs the REFERENCE'S OWN Python test-suite with the oracle riding along. TEST INFRASTRUCTURE.

Build container only (needs /root/reference and oracle/_ref). Every `Simulation` the reference's
tests construct gets an oracle twin built from the same config, map and seed; every `step()` is
replayed on the twin with the same actions and observations, rewards, terminals, truncations and
episode rewards are compared byte for byte; `set_inventory` calls are mirrored. A mismatch raises
inside the reference's test, so it shows up as a failure of that test. Configs the lowering does
not cover (collectives, activation handlers ...) run without a twin and are counted as such.

    python oracle/reference_suite_lockstep.py            # writes tests/golden/reference_suite_lockstep.txt

The summary file is committed so that the pinning travels with the repo; this script is how it was made.
"""

from __future__ import annotations

import os
import sys
import tempfile
from collections import Counter

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.ref_loader import load_reference_python  # noqa: E402

load_reference_python()

import pytest  # noqa: E402
import mettagrid.builder.envs  # noqa: E402,F401  (import order: breaks the config <-> simulator import cycle)
from mettagrid.simulator import simulator as sim_mod  # noqa: E402

from metta_b200.config import MettaGridConfig  # noqa: E402
from metta_b200.lowering import lower_game_config  # noqa: E402
from metta_b200.maps import names_to_codes  # noqa: E402
from oracle.oracle import OracleEnv  # noqa: E402

# A second twin: the CUDA kernel sources themselves, compiled for the host on tests/simt_emu and driven
# through the C-ABI (one env per Simulation). Slower; on by default, KERNEL_TWIN=0 turns it off.
KERNEL_TWIN = os.environ.get("KERNEL_TWIN", "1") != "0"
if KERNEL_TWIN:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from emu_build import load_emu  # noqa: E402
    from harness import HostBuffers  # noqa: E402

    from metta_b200.batch import Batch  # noqa: E402
    from metta_b200.maps import build_pool  # noqa: E402

    _EMU = load_emu()


class _KernelTwin:
    def __init__(self, low, codes, seed):
        self.A = int(low.num_agents)
        self.batch = Batch(low, build_pool(low, codes[None]), 1, clib=_EMU)
        self.batch.set_options(auto_reset=False, map_stride=0)
        self.buf = HostBuffers(self.A, int(low.words[L_OBS_TOKENS]), cuda=False)
        self.batch.set_buffers(*self.buf.ptrs)
        self.batch.reset(np.array([seed], np.uint32), np.array([0], np.int32))

    def step(self, actions):
        self.buf.set_actions(actions)
        self.batch.step()

    def close(self):
        self.batch.close()


from metta_b200.layout import L as _L  # noqa: E402

L_OBS_TOKENS = _L.MGC_NUM_TOKENS

STATS = Counter()
UNCOVERED = Counter()
PER_TEST = {}
_current_test = ["?"]
_last_ctor = {}

_RealCpp = sim_mod.MettaGridCpp


def _capturing_ctor(c_cfg, map_grid, seed):
    _last_ctor["map"], _last_ctor["seed"] = map_grid, seed
    return _RealCpp(c_cfg, map_grid, seed)


sim_mod.MettaGridCpp = _capturing_ctor
_orig_init = sim_mod.Simulation.__init__
_orig_step = sim_mod.Simulation.step
_orig_close = sim_mod.Simulation.close


def _twin_init(self, *args, **kwargs):
    _last_ctor.clear()
    _orig_init(self, *args, **kwargs)
    self._oracle_twin = None
    STATS["simulations"] += 1
    try:
        dump = self._config.model_dump()
        dump["game"]["map_builder"] = None
        cfg = MettaGridConfig.from_reference(dump)
        grid = np.asarray(_last_ctor["map"])
        low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
        twin = OracleEnv(low.words, names_to_codes(low, grid), int(_last_ctor["seed"]))
    except Exception as exc:  # not lowered: outside the hot path's scope
        UNCOVERED[f"{type(exc).__name__}: {str(exc)[:90]}"] += 1
        return
    c = self._Simulation__c_sim
    if not np.array_equal(twin.observations, c.observations()):
        raise AssertionError("oracle twin: initial observations differ from the reference")
    self._oracle_twin = twin
    self._kernel_twin = None
    if KERNEL_TWIN:
        kt = _KernelTwin(low, names_to_codes(low, grid), int(_last_ctor["seed"]))
        if not np.array_equal(kt.buf.obs, c.observations()):
            raise AssertionError("kernel twin: initial observations differ from the reference")
        self._kernel_twin = kt
        STATS["kernel twins"] += 1
    STATS["twinned"] += 1
    PER_TEST[_current_test[0]] = PER_TEST.get(_current_test[0], 0)


def _twin_step(self):
    twin = getattr(self, "_oracle_twin", None)
    c = self._Simulation__c_sim
    actions = np.array(c.actions(), dtype=np.int32).copy() if twin is not None else None
    _orig_step(self)
    if twin is None:
        return
    twin.step(actions)
    for name, got, want in (
        ("observations", twin.observations, c.observations()),
        ("rewards", twin.rewards, c.rewards()),
        ("terminals", twin.terminals, c.terminals()),
        ("truncations", twin.truncations, c.truncations()),
        ("episode rewards", twin.episode_rewards(), c.get_episode_rewards()),
        ("action success", np.asarray(twin.action_success(), dtype=bool), np.asarray(c.action_success(), dtype=bool)),
    ):
        if not np.array_equal(np.asarray(got), np.asarray(want)):
            raise AssertionError(f"oracle twin: {name} differ from the reference at step {c.current_step} ({_current_test[0]})")
    STATS["steps compared"] += 1
    PER_TEST[_current_test[0]] = PER_TEST.get(_current_test[0], 0) + 1
    kt = getattr(self, "_kernel_twin", None)
    if kt is not None:
        kt.step(actions)
        for name, got, want in (
            ("observations", kt.buf.obs, c.observations()),
            ("rewards", kt.buf.rew, c.rewards()),
            ("terminals", kt.buf.term.astype(bool), c.terminals()),
            ("truncations", kt.buf.trunc.astype(bool), c.truncations()),
            ("episode rewards", kt.batch.read_episode_rewards()[0], c.get_episode_rewards()),
            ("action success", kt.batch.read_action_success()[0], np.asarray(c.action_success(), dtype=bool)),
        ):
            if not np.array_equal(np.asarray(got), np.asarray(want)):
                raise AssertionError(f"kernel twin: {name} differ from the reference at step {c.current_step} ({_current_test[0]})")
        STATS["kernel steps compared"] += 1


def _twin_close(self):
    twin = getattr(self, "_oracle_twin", None)
    if twin is not None:
        twin.close()
        self._oracle_twin = None
    kt = getattr(self, "_kernel_twin", None)
    if kt is not None:
        kt.close()
        self._kernel_twin = None
    _orig_close(self)


def _twin_set_inventory(self, inventory):
    ids = {self._sim.resource_names.index(name): int(amount) for name, amount in inventory.items()}
    self._sim._c_sim.set_inventory(self._agent_id, ids)
    twin = getattr(self._sim, "_oracle_twin", None)
    if twin is not None:
        twin.set_inventory(self._agent_id, ids)
        STATS["set_inventory mirrored"] += 1
    kt = getattr(self._sim, "_kernel_twin", None)
    if kt is not None:
        kt.batch.set_inventory(0, self._agent_id, ids)


_orig_grid_objects = sim_mod.Simulation.grid_objects


def _same_objects(ref: dict, mine: dict) -> str | None:
    if set(ref) != set(mine):
        return f"object ids differ: {sorted(set(ref) ^ set(mine))[:6]}"
    for k, ro in ref.items():
        mo = mine[k]
        if set(ro) != set(mo):
            return f"object {k} ({ro.get('type_name')}): keys differ {sorted(set(ro) ^ set(mo))}"
        for f, v in ro.items():
            mv = mo[f]
            if f == "protocols":
                ok = sorted(repr({**x, "inputs": dict(x["inputs"]), "outputs": dict(x["outputs"])}) for x in v) == sorted(map(repr, mv))
            elif f == "vibe_transfers":
                ok = {a: dict(b) for a, b in v.items()} == mv
            elif isinstance(v, dict):
                ok = list(v.items()) == list(mv.items())  # iteration order included (inventory tokens follow it)
            elif isinstance(v, float):
                ok = abs(v - mv) < 1e-6
            else:
                ok = v == mv
            if not ok:
                return f"object {k} ({ro.get('type_name')}): {f} = {mv!r}, reference {v!r}"
    return None


def _twin_grid_objects(self, bbox=None, ignore_types=None):
    ref = _orig_grid_objects(self, bbox, ignore_types)
    kt = getattr(self, "_kernel_twin", None)
    if kt is not None:
        from metta_b200.grid_objects import grid_objects

        if bbox is None:
            mine = grid_objects(kt.batch, 0, ignore_types=tuple(ignore_types or ()))
        else:
            mine = grid_objects(kt.batch, 0, bbox.min_row, bbox.max_row, bbox.min_col, bbox.max_col, tuple(ignore_types or ()))
        why = _same_objects(ref, mine)
        if why is not None:
            raise AssertionError(f"kernel twin: grid_objects() differ from the reference: {why} ({_current_test[0]})")
        STATS["grid_objects compared"] += 1
    return ref


sim_mod.Simulation.grid_objects = _twin_grid_objects
sim_mod.Simulation.__init__ = _twin_init
sim_mod.Simulation.step = _twin_step
sim_mod.Simulation.close = _twin_close
sim_mod.SimulationAgent.set_inventory = _twin_set_inventory


class _Plugin:
    def __init__(self):
        self.outcomes = Counter()
        self.failed = []
        self.uncollected = []

    def pytest_collectreport(self, report):
        if report.failed:
            why = str(report.longrepr).strip().splitlines()[-1] if report.longrepr else ""
            self.uncollected.append((os.path.basename(str(getattr(report, "fspath", "")) or str(report.nodeid)) or "?", why[:120]))

    def pytest_runtest_setup(self, item):
        _current_test[0] = f"{item.path.name}::{item.nodeid.split('::', 1)[-1]}"

    def pytest_runtest_logreport(self, report):
        if report.when == "call" or (report.when == "setup" and report.outcome != "passed"):
            self.outcomes[report.outcome] += 1
            if report.outcome == "failed":
                self.failed.append((f"{os.path.basename(report.location[0])}::{report.nodeid.split('::', 1)[-1]}", str(report.longrepr)[-400:]))


def main():
    tests_dir = "/root/reference/packages/mettagrid/tests"
    if "--cogames" in sys.argv:  # exploratory: the cogames package's tests (not part of the committed summary)
        tests_dir = "/root/reference/packages/cogames/tests"
    files = sorted(f for f in os.listdir(tests_dir) if f.startswith("test_") and f.endswith(".py"))
    plugin = _Plugin()
    with tempfile.TemporaryDirectory() as tmp:
        args = [os.path.join(tests_dir, f) for f in files] + [
            "-q", "-x" if "--stop" in sys.argv else "--continue-on-collection-errors", "-p", "no:cacheprovider",
            "--rootdir", tmp, "-o", "python_files=test_*.py", "--no-header", "-W", "ignore",
        ]
        rc = pytest.main(args, plugins=[plugin])
    covered = {k: v for k, v in PER_TEST.items()}
    lines = [
        "Reference test-suite (packages/mettagrid/tests/test_*.py) run with an oracle twin behind every Simulation",
        "(oracle/reference_suite_lockstep.py, build container, reference core compiled by oracle/build_ref.sh).",
        "",
        f"pytest exit code {int(rc)}; test outcomes: " + ", ".join(f"{k} {v}" for k, v in sorted(plugin.outcomes.items())),
        f"simulations constructed {STATS['simulations']}, with an oracle twin {STATS['twinned']}, "
        f"steps compared byte for byte {STATS['steps compared']}, set_inventory calls mirrored {STATS['set_inventory mirrored']}",
        f"kernel twins (the CUDA sources on tests/simt_emu, through the C-ABI) {STATS['kernel twins']}, "
        f"steps compared byte for byte {STATS['kernel steps compared']}, grid_objects() calls compared {STATS['grid_objects compared']}",
        "",
        "Simulations without a twin (config outside the hot path):",
    ]
    lines += [f"  {n} x {k}" for k, n in UNCOVERED.most_common()] or ["  none"]
    lines += ["", "Failed tests (any 'oracle twin' message is a parity failure; others fail without the twin too):"]
    lines += [f"  {nid}\n      {msg.strip().splitlines()[-1] if msg.strip() else ''}" for nid, msg in plugin.failed] or ["  none"]
    lines += ["", "Test files that could not be imported here (third-party packages absent from this image):"]
    lines += [f"  {n}: {why}" for n, why in plugin.uncollected] or ["  none"]
    lines += ["", "Reference tests that stepped a twinned simulation (test: steps compared):"]
    lines += [f"  {k}: {v}" for k, v in sorted(covered.items()) if v > 0]
    out = os.path.join(ROOT, "tests", "golden", "reference_suite_lockstep.txt")
    if "--cogames" in sys.argv:
        out = "/tmp/reference_suite_lockstep_cogames.txt"
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[:12]))
    print("written", out)


if __name__ == "__main__":
    main()
