"""Builds A/B variants of the library into build_variants/lib_<name>.so (they travel to the GPU box; see
tools/bench_variants.sh). usage: python tools/build_variants.py name=-DFLAG[,-DFLAG2] ..."""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from metta_b200.csrc import build  # noqa: E402

os.makedirs(os.path.join(ROOT, "build_variants"), exist_ok=True)


def one(spec):
    name, _, flags = spec.partition("=")
    out = os.path.join(ROOT, "build_variants", f"lib_{name}.so")
    build.build(force=True, out=out, extra=[f for f in flags.split(",") if f])
    return out


with ThreadPoolExecutor(4) as ex:
    for path in ex.map(one, sys.argv[1:]):
        print(path)
