"""bench.py -- arena agent-steps/s of the batched MettaGrid simulator on B200.

    python bench.py --gpus N --steps K --warmup W            # our arm (torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference core on the host CPUs

Workload (BASELINE.json configs[1]): arena, 4096 envs x 24 agents per GPU, random actions, the
reference's own MapGen maps (256 of them, committed as data), 11x11 window, 200 tokens. One step = one
simulator step of every env.

Steady state: before anything is timed the batch is rolled forward by one whole episode length (10 000
untimed steps) with every env's first episode cut short at its own offset (EarlyResetHandler), so that
the envs sit at evenly staggered phases of their episodes, as they would late in a long training run:
inventories, cooldowns and token counts are those of a running system, and episodes end (auto-reset from
the map pool inside the step kernel) at a rate of N / max_steps per step inside every timed region. The
result therefore does not depend on --steps / --warmup.

Prints ONE JSON line (rank 0). `roofline.frac` is the whole step against the HBM roofline by SURVEY.md
8(d)'s formula (value per GPU x 948 algorithmic bytes / measured copy bandwidth); the two kernels' own
figures are beside it in `roofline.per_kernel`.
"""

from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENVS_PER_GPU = 4096
AGENTS = 24
TOKENS = 200
MAP_POOL = 256
ALG_BYTES_OBS = 890  # per agent-step, observation kernel: 600 obs + 242 window gather + 48 agent record
ALG_BYTES_ACT = 58  # step kernel: 4 action + 4 reward + 2 flags + 48 agent record write-back
ALG_BYTES_STEP = ALG_BYTES_OBS + ALG_BYTES_ACT  # 948, SURVEY.md 8(d)
METRIC = "arena agent-steps/sec"
PHASE_STRIDE = 6181  # env i's first episode is cut at 1 + (i * PHASE_STRIDE) mod (max_steps - 1): low-discrepancy phases, coprime with 9999


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch of the two kernels from the committed ncu capture, with the commit it was taken on."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def cpu_info() -> dict:
    model, phys, cores_per = "unknown", set(), 0
    try:
        pid = None
        for line in open("/proc/cpuinfo"):
            k, _, v = line.partition(":")
            k, v = k.strip(), v.strip()
            if k == "model name":
                model = v
            elif k == "physical id":
                pid = v
                phys.add(v)
            elif k == "cpu cores":
                cores_per = int(v)
    except Exception:
        pass
    return {"model": model, "sockets": len(phys) or None, "physical_cores": (len(phys) or 1) * cores_per or None,
            "logical_cpus": os.cpu_count(), "usable_cpus": len(os.sched_getaffinity(0))}


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ reference / CPU baseline
def _arena_inputs(n_maps):
    from metta_b200.builders import make_arena
    from metta_b200.lowering import lower_game_config
    from metta_b200.maps import arena_mapgen_pool

    cfg = make_arena(AGENTS)
    low = lower_game_config(cfg.game, 62, 62)
    return cfg, low, arena_mapgen_pool(low, n_maps)


def _cpu_worker_loop(conn, kind, envs, first_env, preroll):
    """One host process: `envs` reference (or oracle-port) arena envs, each rolled forward to its own episode
    phase first (the same staggering as the GPU arm), then stepped on request. Replies (agent_steps, seconds)."""
    sys.path.insert(0, ROOT)
    from metta_b200.maps import codes_to_names

    cfg, low, codes = _arena_inputs(MAP_POOL)
    max_steps = int(cfg.game.max_steps)
    rng = np.random.RandomState(1000 + first_env)

    def make(i):
        g = first_env + i
        if kind == "reference":
            from oracle.ref_bridge import RefEnv

            return RefEnv(cfg.game, codes_to_names(low, codes[g % MAP_POOL]), 42 + g, c_cfg=make.c_cfg)
        from oracle.oracle import OracleEnv

        return OracleEnv(low.words, codes[g % MAP_POOL], 42 + g)

    if kind == "reference":
        from oracle.ref_bridge import to_reference_config

        make.c_cfg = to_reference_config(cfg.game)
    sims = [make(i) for i in range(envs)]
    if preroll:
        for i, sim in enumerate(sims):
            phase = ((first_env + i) * PHASE_STRIDE) % (max_steps - 1)
            acts = rng.randint(0, 5, size=(phase, AGENTS)).astype(np.int32)
            for s in range(phase):
                sim.step(acts[s])
    conn.send("ready")
    while True:
        steps = conn.recv()
        if steps is None:
            break
        acts = rng.randint(0, 5, size=(steps, envs, AGENTS)).astype(np.int32)
        t0 = time.perf_counter()
        for s in range(steps):
            for i, sim in enumerate(sims):
                if sim.truncations.all() or sim.terminals.all():  # MettaGridPufferEnv.step: new sim, same seed
                    sims[i] = sim = make(i)
                sim.step(acts[s, i])
        conn.send((envs * AGENTS * steps, time.perf_counter() - t0))


class CpuArm:
    """`procs` worker processes that stay alive between samples (envs are built and pre-rolled once)."""

    def __init__(self, procs: int, envs_per_proc: int, preroll: bool = True):
        from oracle import ref_loader

        self.kind = "reference" if ref_loader.available() else "port"
        if self.kind == "port":
            from oracle import oracle as orc

            orc.build()
        ctx = mp.get_context("fork")
        self.workers = []
        for p in range(procs):
            parent, child = ctx.Pipe()
            pr = ctx.Process(target=_cpu_worker_loop, args=(child, self.kind, envs_per_proc, p * envs_per_proc, preroll), daemon=True)
            pr.start()
            self.workers.append((pr, parent))
        for _, c in self.workers:
            assert c.recv() == "ready"

    def sample(self, steps: int):
        for _, c in self.workers:
            c.send(steps)
        res = [c.recv() for _, c in self.workers]
        slowest = max(r[1] for r in res)
        return sum(r[0] for r in res) / slowest, slowest

    def close(self):
        for pr, c in self.workers:
            c.send(None)
        for pr, _ in self.workers:
            pr.join(timeout=5)


def workload_name(n_envs, agents, tokens):
    return (f"arena, {n_envs} envs x {agents} agents per GPU, random actions, 62x62 maps from the reference's MapGen, "
            f"11x11 window, {tokens} tokens, envs at staggered episode phases (max_steps 10000)")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    envs_per_proc, steps_per_sample = 4, 100
    arm = CpuArm(cores, envs_per_proc)
    times, units = [], 0.0
    for _ in range(args.warmup):
        arm.sample(50)
    for _ in range(args.steps):
        v, secs = arm.sample(steps_per_sample)
        times.append(secs)
        units += v * secs
    arm.close()
    value = units / float(np.sum(times))
    sample = (f"each bench step = {cores} processes x {envs_per_proc} arena envs x {steps_per_sample} env steps of the reference's "
              f"C++ core ({arm.kind}) on the host cores, envs pre-rolled to staggered episode phases; a bounded sample of the "
              f"{args.envs}-env workload")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * float(np.mean(times)), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(args.envs, AGENTS, 200), "envs_per_gpu": args.envs, "agents_per_env": AGENTS, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "agent-steps/s", "cores": cores, "kind": arm.kind, "sample": sample, "cpu": cpu_info()},
        "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------ our arm
class Rig:
    """One batch of envs with torch-owned buffers, a ring of pre-generated actions and an L2 flush."""

    def __init__(self, torch, low, pool, n_envs, agents, tokens, n_actions, dev, seeds, maps, gen_seed, flush_mode="write+read"):
        from metta_b200.batch import Batch

        self.torch, self.dev, self.N, self.A, self.T = torch, dev, n_envs, agents, tokens
        self.batch = Batch(low, pool, n_envs, device=dev.index)
        NA = n_envs * agents
        self.obs = torch.full((NA, tokens, 3), 255, dtype=torch.uint8, device=dev)
        self.term = torch.zeros(NA, dtype=torch.uint8, device=dev)
        self.trunc = torch.zeros(NA, dtype=torch.uint8, device=dev)
        self.rew = torch.zeros(NA, dtype=torch.float32, device=dev)
        # actions: uniformly random and never repeated (a short cycle of pre-generated actions would give every
        # agent a fixed net drift per cycle and pile the agents up against the walls over a long run)
        self.n_actions = n_actions
        self.gen = torch.Generator(device=dev).manual_seed(1234 + gen_seed)
        self.act = torch.zeros(NA, dtype=torch.int32, device=dev)  # the bound action buffer
        self.queue = []  # pre-generated action tensors for timed regions (the policy's output, resident in HBM)
        self.batch.set_buffers(self.obs.data_ptr(), self.term.data_ptr(), self.trunc.data_ptr(), self.rew.data_ptr(), self.act.data_ptr())
        self.batch.reset(seeds, maps)
        self.stream = torch.cuda.current_stream(dev).cuda_stream
        self.flush_mode = flush_mode

    def fresh_actions(self, out=None):
        torch = self.torch
        a = torch.randint(0, self.n_actions, (self.N * self.A,), generator=self.gen, dtype=torch.int32, device=self.dev)
        if out is None:
            return a
        out.copy_(a, non_blocking=True)
        return out

    def prepare(self, k: int):
        """Pre-generate the actions of the next k steps."""
        self.queue = [self.fresh_actions() for _ in range(k)]

    def one_step(self):
        if self.queue:
            self.act.copy_(self.queue.pop(), non_blocking=True)  # what a policy's sampling kernel would write
        else:
            self.fresh_actions(self.act)
        self.batch.step(self.stream)

    def stagger(self, max_steps: int, first_global_env: int):
        """Cut every env's first episode at its own offset and roll the batch forward by one episode length."""
        g = first_global_env + np.arange(self.N, dtype=np.int64)
        self.batch.set_early_reset((1 + (g * PHASE_STRIDE) % (max_steps - 1)).astype(np.uint32))
        for _ in range(max_steps):
            self.one_step()
        self.torch.cuda.synchronize()

    def timed_steps(self, k: int, flush) -> list:
        torch = self.torch
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(k)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(k)]
        self.prepare(k)
        torch.cuda.synchronize()
        for i in range(k):
            flush(i)
            # the step's input is resident before the timed region starts: the actions are written (what a policy's
            # sampling kernel would do) ahead of the start event, the events bracket the env step alone
            self.act.copy_(self.queue.pop(), non_blocking=True)
            starts[i].record()
            self.batch.step(self.stream)
            ends[i].record()
        torch.cuda.synchronize()
        return [s.elapsed_time(e) for s, e in zip(starts, ends)]

    def close(self):
        self.batch.close()


def make_flush(torch, dev, mode):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB of L2
    flush_rd = torch.zeros(64 * 1024 * 1024, dtype=torch.int32, device=dev)  # 256 MiB, only ever read

    def flush_l2(i):
        # write a buffer larger than L2, then stream another one through: the write evicts everything the
        # simulator touched, the read pushes the dirty lines of the write out to HBM so that the timed step
        # does not pay for writing back the flush itself
        flush.fill_(i & 0xFF)
        if mode == "write+read":
            flush_rd.sum()

    return flush_l2


def other_configs(torch, dist, dev, rank, world, flush, peak):
    """Short runs of BASELINE.json configs[2..4]: value, ms per step, whole-step roofline."""
    from metta_b200.builders import make_navigation
    from metta_b200.lowering import lower_game_config
    from metta_b200.maps import build_pool, names_to_codes, random_maps

    out = {}

    def measure(name, rig, steps, note):
        for _ in range(10):
            rig.one_step()
        ms = rig.timed_steps(steps, flush)
        t = torch.tensor([float(sum(ms))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        units = world * rig.N * rig.A * steps
        v = units / (float(t) / 1000.0)
        out[name] = {"value": v, "unit": "agent-steps/s", "ms_per_step": float(t) / steps, "steps": steps,
                     "frac": v / world * ALG_BYTES_STEP / 1e9 / peak, "episodes_finished": float(rig.batch.read_episode_accum()[0]),
                     "workload": note}
        rig.close()

    # C3: navigation / object-use maps with auto-reset, 16384 envs in total, sharded over the ranks
    n3 = 16384 // world
    cfg = make_navigation(num_agents=4, max_steps=100)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 40, 40)
    codes = random_maps(low, 64, 40, 40, {"wall": 120, "assembler": 24, "agent.agent": 4}, seed=3)
    g0 = rank * n3
    rig = Rig(torch, low, build_pool(low, codes), n3, 4, int(cfg.game.obs.num_tokens), len(low.action_names), dev,
              ((np.arange(n3) + g0 + 7) & 0xFFFFFFFF).astype(np.uint32), ((np.arange(n3) + g0) % 64).astype(np.int32), 100 + rank)
    rig.batch.set_options(auto_reset=True, map_stride=5)
    rig.stagger(100, g0)
    measure("navigation_16k", rig, 60, f"configs[2]: navigation/object-use, 40x40 maps (walls, heart assemblers), 4 agents, episodes of 100 steps "
            f"with auto-reset from a 64-map pool, 16384 envs over {world} GPU(s)")
    # C4: a cogames mission, 8192 envs per GPU (the reference's own config and map, from the golden fixture)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from golden_util import load

    cfg, grid, _ = load("cogames_vibe_check")
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    n4, A4 = 8192, cfg.game.num_agents
    rig = Rig(torch, low, build_pool(low, names_to_codes(low, grid)[None]), n4, A4, int(cfg.game.obs.num_tokens), len(low.action_names), dev,
              ((np.arange(n4) + rank * n4 + 11) & 0xFFFFFFFF).astype(np.uint32), np.zeros(n4, np.int32), 200 + rank)
    for _ in range(150):
        rig.one_step()
    measure("cogames_8k_per_gpu", rig, 60, f"configs[3]: cogames VibeCheckMission ({len(low.action_names)} actions uniform, vibe-keyed assemblers, "
            f"chests, chargers, regen), {A4} cogs, 8192 envs per GPU, steps 160+ of the episode")
    return out


def rollout_config(torch, dist, dev, rank, world, peak):
    """configs[4]: env + torch policy in one loop on the device, observations read where the kernel wrote them."""
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    from rollout import run_rollout

    r = run_rollout(envs=ENVS_PER_GPU, steps=100, warmup=20, shim="both", quiet=True)
    v = r["agent_steps_per_s"]
    return {"value": v, "unit": "agent-steps/s", "ms_per_step": r["ms_per_step"], "steps": 100, "frac": v / world * ALG_BYTES_STEP / 1e9 / peak,
            "ms_per_step_by_shim": r["ms_per_step_by_shim"],
            "workload": f"configs[4]: arena rollout loop, {ENVS_PER_GPU} envs x 24 agents per GPU with a bag-of-tokens MLP policy and categorical "
                        f"sampling in the loop, policy input truncated to the kernel's max_flip (mg_set_obs_aux)"}


def host_link_ceiling(torch, dist, dev, world):
    """Plain pinned cudaMemcpyAsync device->host of 10 MB and 60 MB (the sizes of a changed-rows and a full
    export), all ranks at once: what the box's host link gives this rank while its neighbours use it too."""
    out = {}
    for mb in (10, 60):
        d = torch.empty(mb * 1000 * 1000, dtype=torch.uint8, device=dev)
        h = torch.empty(mb * 1000 * 1000, dtype=torch.uint8, pin_memory=True)
        for _ in range(2):
            h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            h.copy_(d, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        gbs = torch.tensor([10 * mb / e0.elapsed_time(e1)], dtype=torch.float64, device=dev)  # MB/ms = GB/s
        lo = gbs.clone()
        if world > 1:
            dist.all_reduce(gbs, op=dist.ReduceOp.SUM)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        out[f"d2h_{mb}MB_gbs_all_ranks"] = float(gbs)
        out[f"d2h_{mb}MB_gbs_slowest_rank"] = float(lo)
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    from metta_b200.maps import build_pool
    from metta_b200.sharding import reduce_episode_accumulators, shard_maps, shard_seeds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    # keep this rank's host threads (and the page-locked buffers they first touch) on the CPU socket its GPU
    # hangs off: with 8 ranks the host-buffer path otherwise crosses sockets
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
    except Exception:
        pass
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    N, A, T = args.envs, AGENTS, TOKENS
    cfg, low, codes = _arena_inputs(MAP_POOL)
    max_steps = int(cfg.game.max_steps)
    pool = build_pool(low, codes)
    # env shards are independent: maps, seeds and episode phases derive from the GLOBAL env index
    seeds = shard_seeds(rank, world, N, 42)
    maps = shard_maps(rank, world, N, MAP_POOL)
    rig = Rig(torch, low, pool, N, A, T, 5, dev, seeds, maps, rank, args.flush)
    batch = rig.batch
    flush_l2 = make_flush(torch, dev, args.flush)
    t_pre = time.perf_counter()
    if not args.no_stagger:
        rig.stagger(max_steps, rank * N)
    preroll_s = time.perf_counter() - t_pre
    steps_now = batch.read_env()["current_step"].astype(np.int64)
    episodes_before = float(batch.read_episode_accum()[0])

    for _ in range(args.warmup):
        rig.one_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = batch.kernel_launches
    # ---- timed region 1: K steps, L2 flushed between steps, CUDA events around each step
    step_ms = rig.timed_steps(args.steps, flush_l2)
    total_ms = float(sum(step_ms))
    launches = batch.kernel_launches - launches0
    episodes_timed = float(batch.read_episode_accum()[0]) - episodes_before
    # ---- free-running: K steps back to back (warm L2, launch gaps included)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rig.prepare(args.steps)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        rig.one_step()
    e1.record()
    torch.cuda.synchronize()
    free_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    # ---- per-kernel durations (whole batch in one launch each, events on the launch stream, L2 flushed)
    act_ms, obs_ms = [], []
    for i in range(min(args.steps, 50)):
        rig.fresh_actions(rig.act)
        flush_l2(i)
        torch.cuda.synchronize()
        a_ms, o_ms = batch.profile_step(rig.stream)
        act_ms.append(a_ms); obs_ms.append(o_ms)
    # ---- end to end through the host-buffer call on the same (staggered) batch: page-locked H2D of the actions,
    # the step, D2H of everything a rollout worker reads
    pin = lambda shape, dtype: torch.empty(shape, dtype=dtype, pin_memory=True).numpy()  # noqa: E731
    e2e_steps = max(3, min(args.steps, 30))
    h_act = pin((2 * (e2e_steps + 3), N * A), torch.int32)  # one fresh row of actions per e2e step
    h_act[...] = np.random.RandomState(rank).randint(0, 5, size=h_act.shape)
    h_obs = pin((N * A, T, 3), torch.uint8); h_rew = pin((N * A,), torch.float32)
    h_term = pin((N * A,), torch.uint8); h_trunc = pin((N * A,), torch.uint8)
    def e2e_run(changed_rows_only):
        batch.set_host_export(changed_rows_only)
        base = (e2e_steps + 3) if changed_rows_only else 0
        for i in range(3):
            batch.step_host(h_act[base + i], h_obs, h_rew, h_term, h_trunc)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        b0 = batch.transfer_bytes()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            batch.step_host(h_act[base + 3 + i], h_obs, h_rew, h_term, h_trunc)
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        b1 = batch.transfer_bytes()
        return secs, (b1[0] - b0[0]) / e2e_steps, (b1[1] - b0[1]) / e2e_steps

    e2e_full_s, _, full_d2h = e2e_run(False)   # every observation byte crosses the host link (mg_set_host_export 0)
    e2e_s, e2e_h2d, e2e_d2h = e2e_run(True)    # only the part of each row that changed (mg_set_host_export 1)
    # the export must leave the host buffer identical to the device observations
    if not np.array_equal(h_obs, batch.read_obs().reshape(h_obs.shape)):
        raise SystemExit("bench: host observation buffer differs from the device buffer after the changed-rows export")
    # ---- the same with the rank's envs split over two handles that are kept in flight together
    # (mg_step_host_begin / _end): one half's observations cross the host link while the other half steps --
    # how the reference's trainer runs its vectorised env (async_factor 2: num_envs = 2 x batch_size,
    # metta/rl/training/training_environment.py:45, batch.py:19)
    halves = []
    for hix in range(2):
        n = N // 2
        sl = slice(hix * n, (hix + 1) * n)
        r = Rig(torch, low, pool, n, A, T, 5, dev, seeds[sl], maps[sl], 100 + 2 * rank + hix, args.flush)
        if not args.no_stagger:
            r.stagger(max_steps, rank * N + hix * n)
        r.batch.set_host_export(True)
        bufs = (pin((n * A, T, 3), torch.uint8), pin((n * A,), torch.float32), pin((n * A,), torch.uint8), pin((n * A,), torch.uint8))
        acts = pin((e2e_steps + 3, n * A), torch.int32)
        acts[...] = np.random.RandomState(1000 + 2 * rank + hix).randint(0, 5, size=acts.shape)
        halves.append((r, bufs, acts))
    for i in range(3):
        for r, bufs, acts in halves:
            r.batch.step_host(acts[i], *bufs)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    hb0 = [r.batch.transfer_bytes() for r, _, _ in halves]
    t0 = time.perf_counter()
    for r, bufs, acts in halves:
        r.batch.step_host_begin(acts[3], *bufs)
    for i in range(1, e2e_steps):
        for r, bufs, acts in halves:
            r.batch.step_host_end()  # this half's step i-1 is on the host: a policy would read it here ...
            r.batch.step_host_begin(acts[3 + i], *bufs)  # ... and answer with the next actions, while the other half is in flight
    for r, _, _ in halves:
        r.batch.step_host_end()
    torch.cuda.synchronize()
    e2e_halves_s = time.perf_counter() - t0
    hb1 = [r.batch.transfer_bytes() for r, _, _ in halves]
    halves_h2d = sum(b1[0] - b0[0] for b0, b1 in zip(hb0, hb1)) / e2e_steps
    halves_d2h = sum(b1[1] - b0[1] for b0, b1 in zip(hb0, hb1)) / e2e_steps
    for r, bufs, _ in halves:
        if not np.array_equal(bufs[0], r.batch.read_obs().reshape(bufs[0].shape)):
            raise SystemExit("bench: host observation buffer differs from the device buffer after the two-halves run")
        r.close()
    del halves
    link = host_link_ceiling(torch, dist, dev, world)
    # ---- the only collective of the env path: episode stats (a few KB) over NCCL
    acc = torch.from_numpy(batch.read_episode_accum()).to(dev)
    t = torch.tensor([total_ms, free_ms, e2e_s, e2e_full_s, e2e_halves_s], dtype=torch.float64, device=dev)
    per_rank = torch.zeros(world, dtype=torch.float64, device=dev)
    per_rank[rank] = total_ms / args.steps
    ep = torch.tensor([episodes_timed], dtype=torch.float64, device=dev)
    reduce_episode_accumulators(acc)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(per_rank, op=dist.ReduceOp.SUM)
        dist.all_reduce(ep, op=dist.ReduceOp.SUM)
    total_ms, free_ms, e2e_s, e2e_full_s, e2e_halves_s = (float(x) for x in t.cpu())
    peak, peak_kind = peaks()
    rig.close()
    others = None
    if not args.no_other_configs:
        others = other_configs(torch, dist, dev, rank, world, flush_l2, peak)
        others["rollout_arena"] = rollout_config(torch, dist, dev, rank, world, peak)
    if rank == 0:
        units = world * N * A * args.steps
        value = units / (total_ms / 1000.0)
        obs_mean_ms = float(np.mean(obs_ms)); act_mean_ms = float(np.mean(act_ms))
        achieved = value / world * ALG_BYTES_STEP / 1e9
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(1, 8)
            arm.sample(100)
            v, secs = arm.sample(12000)
            arm.close()
            cpu = {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": arm.kind, "cpu": cpu_info(),
                   "sample": f"8 arena envs x 24 agents x 12000 steps on one host core, envs pre-rolled to staggered episode phases ({secs:.1f} s)"}
        traffic = ncu_traffic() if N == ENVS_PER_GPU else None
        e2e_value = world * N * A * e2e_steps / e2e_s
        e2e_full_value = world * N * A * e2e_steps / e2e_full_s
        d2h_rate = e2e_d2h * e2e_steps / e2e_s / 1e9  # GB/s this rank moved in changed-rows mode
        e2e_halves_value = world * 2 * (N // 2) * A * e2e_steps / e2e_halves_s
        halves_rate = halves_d2h * e2e_steps / e2e_halves_s / 1e9
        full_rate = full_d2h * e2e_steps / e2e_full_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(N, A, T),
                       "envs_per_gpu": N, "agents_per_env": A, "l2": "flushed between timed steps (256 MiB fill" + (", then 256 MiB streamed read)" if args.flush == "write+read" else ")"),
                       "parallelism": f"env shards x{world}, no data-path collective",
                       "timed_region": "CUDA events around mg_step alone; each step's actions (pre-generated, resident in HBM) are "
                                       "copied into the bound action buffer before the start event, after the L2 flush "
                                       "(free_running has that copy inside and a warm L2)",
                       "steady_state": {"preroll_steps": 0 if args.no_stagger else max_steps, "preroll_s": preroll_s,
                                        "episode_step_min": int(steps_now.min()), "episode_step_median": int(np.median(steps_now)),
                                        "episode_step_max": int(steps_now.max())}},
            "clocks": clocks,
            "per_rank_ms": [float(x) for x in per_rank.cpu()],
            "step_ms": {"min": float(np.min(step_ms)), "median": float(np.median(step_ms)), "p90": float(np.percentile(step_ms, 90)), "max": float(np.max(step_ms))},
            "e2e": {"value": e2e_halves_value, "unit": "agent-steps/s",
                    "h2d_bytes_per_step": int(halves_h2d), "d2h_bytes_per_step": int(halves_d2h),
                    "mode": "two_halves: the rank's envs as two handles of N/2 envs kept in flight together (mg_step_host_begin / mg_step_host_end, "
                            "the reference trainer's async_factor 2), page-locked caller buffers, mg_set_host_export 1 (observations as changed row "
                            "pieces written by a kernel); every step's actions go host->device and its observations, rewards and flags come back "
                            "inside the timed region; host buffers checked equal to the device buffers after the run. The single-handle numbers "
                            "of the same call are under modes",
                    "modes": {"two_halves": {"value": e2e_halves_value, "d2h_bytes_per_step": int(halves_d2h), "d2h_gbs_per_rank": halves_rate},
                              "changed_rows": {"value": e2e_value, "d2h_bytes_per_step": int(e2e_d2h), "d2h_gbs_per_rank": d2h_rate,
                                               "note": "one handle, mg_step_host with mg_set_host_export 1 (what GpuMettaGridVecEnv(host=True) uses)"},
                              "full_rows": {"value": e2e_full_value, "d2h_bytes_per_step": int(full_d2h), "d2h_gbs_per_rank": full_rate,
                                            "note": "same call with every observation byte copied (mg_set_host_export 0, the C-ABI default)"}},
                    "host_link": link,
                    "host_link_frac": {"two_halves": halves_rate * world / link["d2h_10MB_gbs_all_ranks"],
                                       "changed_rows": d2h_rate * world / link["d2h_10MB_gbs_all_ranks"], "full_rows": full_rate * world / link["d2h_60MB_gbs_all_ranks"]}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "whole step: mg_step_kernel + mg_obs_kernel in overlapped slices", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "alg_bytes_per_agent_step": ALG_BYTES_STEP,
                         "formula": "value per GPU x 948 B / peak (SURVEY.md 8d)", "peak_source": peak_kind,
                         "traffic": (traffic or {}).get("whole_step_bytes"), "traffic_source": (traffic or {}).get("source"),
                         "per_kernel": {
                             "mg_step_kernel": {"ms": act_mean_ms, "alg_bytes_per_launch": ALG_BYTES_ACT * N * A,
                                                "achieved": ALG_BYTES_ACT * N * A / (act_mean_ms / 1000.0) / 1e9,
                                                "frac": ALG_BYTES_ACT * N * A / (act_mean_ms / 1000.0) / 1e9 / peak,
                                                "traffic": (traffic or {}).get("mg_step_kernel_bytes")},
                             "mg_obs_kernel": {"ms": obs_mean_ms, "alg_bytes_per_launch": ALG_BYTES_OBS * N * A,
                                               "achieved": ALG_BYTES_OBS * N * A / (obs_mean_ms / 1000.0) / 1e9,
                                               "frac": ALG_BYTES_OBS * N * A / (obs_mean_ms / 1000.0) / 1e9 / peak,
                                               "traffic": (traffic or {}).get("mg_obs_kernel_bytes")},
                             "note": "each kernel alone over the whole batch (mg_profile_step), L2 flushed; inside a step they run as slices that overlap"}},
            "free_running": {"value": world * N * A * args.steps / (free_ms / 1000.0), "unit": "agent-steps/s", "l2": "warm"},
            "cpu_baseline": cpu,
            "episodes_finished": float(ep),
            "episodes_finished_since_start": float(acc[0]),
            "other_configs": others,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    ap.add_argument("--no-stagger", action="store_true", help="skip the pre-roll: time the first steps of fresh episodes (tuning only)")
    ap.add_argument("--flush", default="write+read", choices=["write", "write+read"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        args.steps = min(args.steps, 200)  # each step is a bounded CPU sample (about half a second)
        args.warmup = min(args.warmup, 5)
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
