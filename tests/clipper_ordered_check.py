"""Run by tests/test_oracle_vs_reference.py in a process of its own with MG_REF_DIR=_ref_ordered: the oracle against
the reference build whose clipper sets are ordered by object id (oracle/build_ref_ordered.sh). TEST INFRASTRUCTURE."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from harness import random_actions  # noqa: E402
from scenarios import SCENARIOS  # noqa: E402

from metta_b200.lowering import lower_game_config  # noqa: E402
from metta_b200.maps import names_to_codes  # noqa: E402
from oracle.oracle import OracleEnv  # noqa: E402
from oracle.ref_bridge import RefEnv  # noqa: E402

assert os.environ.get("MG_REF_DIR") == "_ref_ordered"
cfg, grid = SCENARIOS["clipper_spread"]()
low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
codes = names_to_codes(low, grid)
for seed in range(30, 36):
    ref = RefEnv(cfg.game, grid.tolist(), seed)
    orc = OracleEnv(low.words, codes, seed)
    ids = sorted(k for k, o in ref.grid_objects().items() if "is_clipped" in o)
    rng = np.random.RandomState(seed)
    changes, prev = 0, None
    for s in range(1200):
        a = random_actions(rng, cfg, low, cfg.game.num_agents)
        ref.step(a)
        orc.step(a)
        assert np.array_equal(ref.observations, orc.observations) and np.array_equal(ref.rewards, orc.rewards), (seed, s)
        if s % 5 == 0:
            objs = ref.grid_objects()
            rc = [int(bool(objs[k]["is_clipped"])) for k in ids]
            assert rc == [int(r[2]) for r in orc.assembler_state()], (seed, s)
            changes += prev is not None and rc != prev
            prev = rc
    assert changes > 50, changes
print("ok")
