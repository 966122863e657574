"""The kernel sources on the SIMT emulator once more, this time built with AddressSanitizer: 'device'
buffers are host allocations there, so an out-of-bounds access of a kernel shows up as an ASan report.
(compute-sanitizer is not available on the GPU pool.) Runs in a subprocess because ASan has to be
loaded first."""

import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _libasan():
    try:
        p = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    except OSError:
        return None
    return p if os.path.sep in p and os.path.exists(p) else None


@pytest.mark.skipif(_libasan() is None, reason="libasan not installed")
def test_emulated_kernels_are_asan_clean(tmp_path):
    so = tmp_path / "libmg_emu_asan.so"
    csrc = os.path.join(ROOT, "metta_b200", "csrc")
    emu = os.path.join(ROOT, "tests", "simt_emu")
    subprocess.check_call([
        "g++", "-std=c++17", "-O1", "-g", "-fPIC", "-shared", "-DMG_SIMT_EMU", "-ffp-contract=off", "-fsanitize=address",
        "-fno-omit-frame-pointer", "-I", emu, "-I", csrc, "-x", "c++", os.path.join(csrc, "mg_abi.cu"),
        os.path.join(csrc, "mg_host_order.cpp"), os.path.join(emu, "simt_emu.cpp"), "-o", str(so),
    ])  # fmt: skip
    script = textwrap.dedent(f"""
        import ctypes, sys
        sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})
        import numpy as np
        from metta_b200 import capi
        lib = capi._declare(ctypes.CDLL({str(so)!r}))
        from harness import Lockstep, random_actions
        from scenarios import SCENARIOS
        from golden_util import load
        from metta_b200.lowering import lower_game_config
        from metta_b200.maps import names_to_codes, arena_maps
        from metta_b200.builders import make_arena
        for name, steps in (("kitchen_sink", 60), ("tiny_truncating", 17), ("clipped_start", 40)):
            cfg, grid = SCENARIOS[name]()
            low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
            ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=3, seeds=[3, 4, 5], clib=lib)
            rng = np.random.RandomState(11)
            for s in range(steps):
                ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=29, step=s))
            ls.close()
        cfg, grid, _ = load("cogames_vibe_check")
        low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
        ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=2, seeds=[5, 6], clib=lib)
        rng = np.random.RandomState(21)
        for s in range(30):
            ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
        ls.close()
        cfg = make_arena(24)
        low = lower_game_config(cfg.game, 62, 62)
        ls = Lockstep(cfg, arena_maps(low, 2, seed=9), num_envs=2, seeds=[42, 43], clib=lib)
        rng = np.random.RandomState(2)
        for s in range(40):
            ls.step(rng.randint(0, 5, size=ls.N * ls.A))
        ls.close()
        print("ASAN-CLEAN")
    """)
    env = dict(os.environ, LD_PRELOAD=_libasan(), ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0")
    r = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, env=env, timeout=1500)
    assert "ASAN-CLEAN" in r.stdout and "ERROR: AddressSanitizer" not in r.stderr, r.stdout[-1500:] + r.stderr[-3000:]
