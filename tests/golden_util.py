import json
import os

import numpy as np

from metta_b200.config import MettaGridConfig

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = ["arena", "kitchen_sink", "tiny_truncating", "clipped_start", "cogames_vibe_check", "cogames_machina_chests",
         # clipper picks, from the reference build with id-ordered clipper sets (oracle/build_ref_ordered.sh)
         "clipper_spread", "cogames_clip_period"]


def load(name: str):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        cfg = MettaGridConfig.from_reference(json.load(f))
    data = np.load(os.path.join(GOLDEN, name + ".npz"))
    return cfg, data["grid"], data
