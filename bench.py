"""bench.py -- arena agent-steps/s of the batched MettaGrid simulator on B200.

    python bench.py --gpus N --steps K --warmup W            # our arm (torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference core on the host CPUs

Workload (BASELINE.json configs[1]): arena, 4096 envs x 24 agents per GPU, random actions,
synthetic maps, 11x11 window, 200 tokens. One step = one simulator step of every env.
Prints ONE JSON line (rank 0).
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
ALG_BYTES_ACT = 58  # action kernel: 4 action + 4 reward + 2 flags + 48 agent record write-back
# DRAM traffic of one observation-kernel launch over 4096 envs x 24 agents, from the committed ncu capture
# (profiles/r01k_ncu_full_summary.txt): 68.92 MB read + 23.40 MB written (algorithmic: 87.49 MB)
NCU_OBS_TRAFFIC_BYTES = 68916736 + 23396352
METRIC = "arena agent-steps/sec"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


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
def _cpu_worker(args):
    """Steps `envs` reference (or oracle-port) envs for `steps` steps; returns (agent_steps, seconds)."""
    kind, envs, steps, seed = args
    sys.path.insert(0, ROOT)
    from metta_b200.builders import make_arena
    from metta_b200.lowering import lower_game_config
    from metta_b200.maps import arena_maps, codes_to_names

    cfg = make_arena(AGENTS)
    low = lower_game_config(cfg.game, 62, 62)
    codes = arena_maps(low, envs, seed=seed)
    if kind == "reference":
        from oracle.ref_bridge import RefEnv, to_reference_config

        c_cfg = to_reference_config(cfg.game)
        sims = [RefEnv(cfg.game, codes_to_names(low, codes[i]), seed + i, c_cfg=c_cfg) for i in range(envs)]
    else:
        from oracle.oracle import OracleEnv

        sims = [OracleEnv(low.words, codes[i], seed + i) for i in range(envs)]
    rng = np.random.RandomState(seed)
    acts = rng.randint(0, 5, size=(steps, envs, AGENTS)).astype(np.int32)
    for s in range(min(20, steps)):  # warm-up
        for i, sim in enumerate(sims):
            sim.step(acts[s, i])
    t0 = time.perf_counter()
    for s in range(steps):
        for i, sim in enumerate(sims):
            sim.step(acts[s, i])
    dt = time.perf_counter() - t0
    return envs * AGENTS * steps, dt


def cpu_throughput(procs: int, envs_per_proc: int, steps: int):
    from oracle import ref_loader

    kind = "reference" if ref_loader.available() else "port"
    if kind == "port":
        from oracle import oracle as orc

        orc.build()
    jobs = [(kind, envs_per_proc, steps, 1000 + 17 * p) for p in range(procs)]
    if procs == 1:
        res = [_cpu_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            res = pool.map(_cpu_worker, jobs)
    total = sum(r[0] for r in res)
    slowest = max(r[1] for r in res)
    return total / slowest, kind, slowest


def workload_name(n_envs, agents, tokens):
    return f"arena, {n_envs} envs x {agents} agents per GPU, random actions, 62x62 synthetic maps, 11x11 window, {tokens} tokens"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    envs_per_proc, steps_per_sample = 4, 100
    times, units, kind = [], 0.0, "port"
    for _ in range(args.warmup):
        cpu_throughput(cores, envs_per_proc, 50)
    for _ in range(args.steps):
        v, kind, secs = cpu_throughput(cores, envs_per_proc, steps_per_sample)
        times.append(secs)
        units += v * secs
    value = units / float(np.sum(times))
    sample = (f"each bench step = {cores} processes x {envs_per_proc} arena envs x {steps_per_sample} env steps of the reference's "
              f"C++ core ({kind}) on the host cores, a bounded sample of the {args.envs}-env workload")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * float(np.mean(times)), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(args.envs, AGENTS, 200), "envs_per_gpu": args.envs, "agents_per_env": AGENTS, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "agent-steps/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    from metta_b200.batch import Batch
    from metta_b200.builders import make_arena
    from metta_b200.lowering import lower_game_config
    from metta_b200.maps import arena_maps, build_pool
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
    cfg = make_arena(A)
    low = lower_game_config(cfg.game, 62, 62)
    # env shards are independent: maps and seeds derive from the GLOBAL env index
    codes = arena_maps(low, MAP_POOL, seed=1234)
    pool = build_pool(low, codes)
    batch = Batch(low, pool, N, device=local)
    obs = torch.full((N * A, T, 3), 255, dtype=torch.uint8, device=dev)
    term = torch.zeros(N * A, dtype=torch.uint8, device=dev)
    trunc = torch.zeros(N * A, dtype=torch.uint8, device=dev)
    rew = torch.zeros(N * A, dtype=torch.float32, device=dev)
    g = torch.Generator().manual_seed(rank)
    ring = 16
    act_ring = torch.randint(0, 5, (ring, N * A), generator=g, dtype=torch.int32).to(dev)
    act = act_ring[0].clone()  # the bound action buffer; every step starts by refreshing it from the ring (device copy)
    batch.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
    seeds = shard_seeds(rank, world, N, 42)
    maps = shard_maps(rank, world, N, MAP_POOL)
    batch.reset(seeds, maps)
    stream = torch.cuda.current_stream(dev).cuda_stream
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB of L2
    flush_rd = torch.zeros(64 * 1024 * 1024, dtype=torch.int32, device=dev)  # 256 MiB, only ever read

    def flush_l2(i):
        # write a buffer larger than L2, then stream another one through: the write evicts everything
        # the simulator touched, the read pushes the dirty lines of the write out to HBM so that the
        # timed step does not pay for writing back the flush itself
        flush.fill_(i & 0xFF)
        if args.flush == "write+read":
            flush_rd.sum()

    def one_step(i):
        act.copy_(act_ring[i % ring], non_blocking=True)  # what a policy's sampling kernel would write
        batch.step(stream)

    for i in range(args.warmup):
        one_step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = batch.kernel_launches
    # ---- timed region 1: K steps, L2 flushed between steps, CUDA events around each step
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    torch.cuda.synchronize()
    for i in range(args.steps):
        flush_l2(i)
        starts[i].record()
        one_step(args.warmup + i)
        ends[i].record()
    torch.cuda.synchronize()
    step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    total_ms = float(sum(step_ms))
    launches = batch.kernel_launches - launches0
    # ---- free-running: K steps back to back (warm L2, launch gaps included)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        one_step(i)
    e1.record()
    torch.cuda.synchronize()
    free_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    # ---- per-kernel durations for the roofline (events on the launch stream, L2 flushed)
    act_ms, obs_ms = [], []
    for i in range(min(args.steps, 50)):
        act.copy_(act_ring[i % ring])
        flush_l2(i)
        torch.cuda.synchronize()
        a_ms, o_ms = batch.profile_step(stream)
        act_ms.append(a_ms); obs_ms.append(o_ms)
    # ---- end to end through the host-buffer call: pinned H2D of actions, D2H of everything
    hb = Batch(low, pool, N, device=local)
    hb.use_internal_buffers()
    hb.reset(seeds, maps)
    # page-locked host buffers (what a host-side rollout worker would hold): the copy engines read/write them directly
    pin = lambda shape, dtype: torch.empty(shape, dtype=dtype, pin_memory=True).numpy()  # noqa: E731
    h_act = pin((ring, N * A), torch.int32)
    h_act[...] = np.random.RandomState(rank).randint(0, 5, size=(ring, N * A))
    h_obs = pin((N * A, T, 3), torch.uint8); h_rew = pin((N * A,), torch.float32)
    h_term = pin((N * A,), torch.uint8); h_trunc = pin((N * A,), torch.uint8)
    e2e_steps = max(3, min(args.steps, 30))

    def e2e_run(changed_rows_only):
        hb.set_host_export(changed_rows_only)
        for i in range(3):
            hb.step_host(h_act[i % ring], h_obs, h_rew, h_term, h_trunc)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        b0 = hb.transfer_bytes()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            hb.step_host(h_act[i % ring], h_obs, h_rew, h_term, h_trunc)
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        b1 = hb.transfer_bytes()
        return secs, (b1[0] - b0[0]) / e2e_steps, (b1[1] - b0[1]) / e2e_steps

    e2e_full_s, _, full_d2h = e2e_run(False)   # every observation byte crosses the host link
    e2e_s, e2e_h2d, e2e_d2h = e2e_run(True)    # only the part of each row that changed (mg_set_host_export 1)
    # the export must leave the host buffer identical to the device observations
    if not np.array_equal(h_obs, hb.read_obs().reshape(h_obs.shape)):
        raise SystemExit("bench: host observation buffer differs from the device buffer after the changed-rows export")
    # ---- the only collective of the env path: episode stats (a few KB) over NCCL
    acc = torch.from_numpy(batch.read_episode_accum()).to(dev)
    t = torch.tensor([total_ms, free_ms, e2e_s, e2e_full_s], dtype=torch.float64, device=dev)
    reduce_episode_accumulators(acc)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, free_ms, e2e_s, e2e_full_s = (float(x) for x in t.cpu())
    if rank == 0:
        units = world * N * A * args.steps
        value = units / (total_ms / 1000.0)
        peak, peak_kind = peaks()
        obs_mean_ms = float(np.mean(obs_ms)); act_mean_ms = float(np.mean(act_ms))
        achieved = ALG_BYTES_OBS * N * A / (obs_mean_ms / 1000.0) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, kind, secs = cpu_throughput(1, 16, 9000)
            cpu = {"value": v, "unit": "agent-steps/s", "cores": 1, "kind": kind,
                   "sample": f"16 arena envs x 24 agents x 9000 steps on one host core ({secs:.1f} s)"}
        line = {
            "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(N, A, T),
                       "envs_per_gpu": N, "agents_per_env": A, "l2": "flushed between timed steps (256 MiB fill" + (", then 256 MiB streamed read)" if args.flush == "write+read" else ")"),
                       "parallelism": f"env shards x{world}, no data-path collective"},
            "clocks": clocks,
            "e2e": {"value": world * N * A * e2e_steps / e2e_s, "unit": "agent-steps/s",
                    "h2d_bytes_per_step": int(e2e_h2d), "d2h_bytes_per_step": int(e2e_d2h),
                    "host_export": "page-locked caller buffers; observations: changed row prefixes written by a kernel (mg_set_host_export 1), "
                                   "host buffer checked equal to the device buffer after the run",
                    "full_rows": {"value": world * N * A * e2e_steps / e2e_full_s, "d2h_bytes_per_step": int(full_d2h),
                                  "note": "same call with every observation byte copied (mg_set_host_export 0)"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "mg_obs_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": NCU_OBS_TRAFFIC_BYTES if N == ENVS_PER_GPU else None,
                         "traffic_source": "profiles/r01k_ncu_full_summary.txt (ncu --set full, one launch over 4096 envs: dram__bytes_read.sum + dram__bytes_write.sum)",
                         "peak_source": peak_kind,
                         "kernel_ms": obs_mean_ms, "alg_bytes_per_launch": ALG_BYTES_OBS * N * A,
                         "step_kernel_ms": act_mean_ms, "whole_step_frac": (ALG_BYTES_OBS + ALG_BYTES_ACT) * N * A / ((obs_mean_ms + act_mean_ms) / 1000.0) / 1e9 / peak},
            "free_running": {"value": world * N * A * args.steps / (free_ms / 1000.0), "unit": "agent-steps/s", "l2": "warm"},
            "cpu_baseline": cpu,
            "episodes_finished": float(acc[0]),
        }
        print(json.dumps(line))
    hb.close()
    batch.close()
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
