"""mg_peek_observations_noalloc: the entry point for callers that must not allocate GPU memory.

A monitor or replay thread that looks at a running rollout cannot always call cudaMalloc: the
trainer's caching allocator may hold nearly all device memory, and cudaMalloc/cudaFree synchronise
the whole device. This entry point therefore keeps the observation kernel's scratch (per-agent
token counts) in a fixed 1 MiB device array inside the library and writes into a buffer the caller
already owns. The tests call it through the C-ABI and compare with the regular observation pass
(mg_compute_observations), twice in a row with the state changing in between, and check that
batches beyond the fixed capacity are refused with a clear message.

This file sorts last on purpose: if it fails, `pytest -x` has already run everything else.
"""

import numpy as np
import pytest
from scenarios import SCENARIOS

from metta_b200 import capi
from metta_b200.batch import ST_OBS, Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_maps, build_pool, names_to_codes
from harness import random_actions


def _setup(which):
    if which == "arena":
        cfg = make_arena(24)
        low = lower_game_config(cfg.game, 62, 62)
        codes = arena_maps(low, 4, seed=21)
        n = 12
    else:  # 48 tokens per agent: observations overflow, the count scratch sees attempted > written
        cfg, grid = SCENARIOS["kitchen_sink"]()
        low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
        codes = names_to_codes(low, grid)[None]
        n = 9
    return cfg, low, build_pool(low, codes), n


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["arena", "kitchen_sink"])
def test_peek_matches_regular_observation_pass(which, product_lib):
    import torch

    assert torch.cuda.is_available()
    cfg, low, pool, n = _setup(which)
    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, pool, n)
    obs = torch.zeros((n * A, T, 3), dtype=torch.uint8, device="cuda")
    term = torch.zeros(n * A, dtype=torch.uint8, device="cuda")
    trunc = torch.zeros(n * A, dtype=torch.uint8, device="cuda")
    rew = torch.zeros(n * A, dtype=torch.float32, device="cuda")
    act = torch.zeros(n * A, dtype=torch.int32, device="cuda")
    b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
    b.reset(np.arange(n, dtype=np.uint32) + 5, (np.arange(n) % len(pool.codes)).astype(np.int32))
    side = torch.full_like(obs, 7)  # the caller's own buffer; allocated before the calls under test
    rng = np.random.RandomState(8)
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):  # two calls in a row, the state moves on in between
        for s in range(6):
            act.copy_(torch.from_numpy(random_actions(rng, cfg, low, n * A)))
            b.step(stream)
        b.compute_observations(stream)  # the existing entry point, into the bound buffer
        torch.cuda.synchronize()
        expected = obs.clone()
        stats_before = b.read_game_stats().copy()
        b.peek_observations(side.data_ptr(), stream)  # the allocation-free entry point
        torch.cuda.synchronize()
        assert torch.equal(side, expected), f"call {call}: peeked observations differ from the regular pass"
        assert torch.equal(obs, expected), "the bound buffer must not be touched"
        assert np.array_equal(b.read_game_stats(), stats_before), "token statistics must not be touched"
    b.close()


@pytest.mark.gpu
def test_peek_rejects_batches_beyond_the_static_scratch(product_lib):
    import torch

    cfg, grid = SCENARIOS["tiny_truncating"]()  # 2 agents per env
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    pool = build_pool(low, names_to_codes(low, grid)[None])
    n = 262144 // 2 + 1
    b = Batch(low, pool, n)
    b.use_internal_buffers()
    side = torch.zeros((n * 2, cfg.game.obs.num_tokens, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(capi.MgError, match="exceed the static scratch of 262144 agents"):
        b.peek_observations(side.data_ptr())
    with pytest.raises(capi.MgError, match="null handle or output buffer"):
        b.peek_observations(0)
    b.close()


def test_peek_logic_on_the_emulated_kernels(emu_lib):
    """Same comparison on the CPU build of the kernel sources (no GPU): output equality, twice."""
    cfg, low, pool, n = _setup("kitchen_sink")
    n = 2
    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    b = Batch(low, pool, n, clib=emu_lib)
    obs = np.zeros((n * A, T, 3), np.uint8); term = np.zeros(n * A, np.uint8); trunc = np.zeros(n * A, np.uint8)
    rew = np.zeros(n * A, np.float32); act = np.zeros(n * A, np.int32)
    b.set_buffers(obs.ctypes.data, term.ctypes.data, trunc.ctypes.data, rew.ctypes.data, act.ctypes.data)
    b.reset(np.array([3, 4], np.uint32), np.zeros(n, np.int32))
    side = np.full_like(obs, 7)
    rng = np.random.RandomState(1)
    for call in range(2):
        for s in range(5):
            act[:] = random_actions(rng, cfg, low, n * A)
            b.step()
        b.compute_observations()
        expected = obs.copy()
        b.peek_observations(side.ctypes.data)
        assert np.array_equal(side, expected) and np.array_equal(obs, expected)
    b.close()
