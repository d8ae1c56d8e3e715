#!/usr/bin/env python
"""bench.py -- SALP env-steps/s on N B200s (one process per GPU), roofline, CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A "step" is ONE launch of the env-step kernel over this rank's slab of environments (one env
step for every env).  Workloads (BASELINE.json configs; per-GPU slab, weak scaling -- envs never
interact, so ranks share nothing on the data path):
    single_food_1m   configs/single_food.yaml, 1,048,576 envs / GPU, fp32          [default]
    long_horizon_64k configs/single_food_long_horizon.yaml, 65,536 envs / GPU, fp32 (configs[2])
    defaults_1m      configs/defaults.yaml (12 food), 1,048,576 envs / GPU, fp32   (configs[3])
    sac_rollout_256k configs/sac_gail.yaml, 262,144 envs / GPU, actor MLP + env step + replay storage (configs[4])
    parity_4k        configs/single_food.yaml, 4,096 envs, fp64 validation build   (configs[1])
Prints ONE JSON line (rank 0).  See DESIGN.md section "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SINGLE_FOOD = dict(num_food_items=1, food_reward=1000.0, collision_penalty=-10.0, time_penalty=-0.1,
                   proximity_reward_weight=5.0, respawn_food=True, forced_breathing=True,
                   max_steps_without_food=1500, efficiency_bonus=0.0)
LONG_HORIZON = dict(num_food_items=1, food_reward=1000.0, collision_penalty=-500.0, time_penalty=-0.5,
                    proximity_reward_weight=3.0, respawn_food=True, forced_breathing=True,
                    max_steps_without_food=2000, efficiency_bonus=0.0)
DEFAULTS = dict(num_food_items=12, food_reward=15.0, collision_penalty=-30.0, time_penalty=-0.05,
                efficiency_bonus=2.0, forced_breathing=True)
WORKLOADS = {
    "single_food_1m": dict(yaml="configs/single_food.yaml", params=SINGLE_FOOD, envs=1 << 20, dtype="float32"),
    "long_horizon_64k": dict(yaml="configs/single_food_long_horizon.yaml", params=LONG_HORIZON, envs=1 << 16, dtype="float32"),
    "defaults_1m": dict(yaml="configs/defaults.yaml", params=DEFAULTS, envs=1 << 20, dtype="float32"),
    "parity_4k": dict(yaml="configs/single_food.yaml", params=SINGLE_FOOD, envs=1 << 12, dtype="float64"),
    # configs[4]: SAC rollout collection -- [24->256->256->2] tanh-Gaussian actor (torch, bf16) -> env kernel ->
    # replay-append kernel, replayed from a CUDA graph; a "step" is one env step of that loop for every env
    "sac_rollout_256k": dict(yaml="configs/sac_gail.yaml", params=DEFAULTS, envs=1 << 18, dtype="float32", rollout=True),
    # model (B), one env step = one breathing cycle (0..1450 integrator substeps, 725 on average
    # for actions uniform in the env's Box); ALU-bound, reported beside its (tiny) HBM figure
    "robot_cycle_256k": dict(yaml=None, params={}, envs=1 << 18, dtype="float32", cycle=True),
}
RING_STEPS = 5  # rows of the zero-copy replay ring: (5 - 1) x 262144 envs = 1 048 576 transitions (the reference's buffer_size)
METRIC = "salp_env_steps_per_sec"
UNIT = "env-steps/s"


def algorithmic_bytes_per_env_step(n_food: int, real_bytes: int, k_fused: int = 1, act_dim: int = 1, obs_dim: int = 24) -> float:
    """B(K) = A + O + R + D + (S_in + S_out)/K,
    S = 8 reals + 4 (fsm) + 8 (counters) + 8 (cursors) + (real + 4) (ep return, len) + 2F reals + 4 (mask)."""
    A, O, R, D = real_bytes * act_dim, 4 * obs_dim, real_bytes, 2
    S = 8 * real_bytes + 4 + 8 + 8 + (real_bytes + 4) + 2 * n_food * real_bytes + 4
    return A + O + R + D + 2.0 * S / k_fused


def min_traffic_bytes_per_env_step(n_food: int, real_bytes: int, k_fused: int = 1, act_dim: int = 1, obs_dim: int = 24) -> float:
    """Tighter count than SURVEY 8(d): food records are read every launch but only rewritten on a respawn
    (~1/1000 steps), so the state write-back is S minus the 2F food reals; events (1 B) are written too."""
    A, O, R, D = real_bytes * act_dim, 4 * obs_dim, real_bytes, 3
    S = 8 * real_bytes + 4 + 8 + 8 + (real_bytes + 4) + 2 * n_food * real_bytes + 4
    return A + O + R + D + (2.0 * S - 2 * n_food * real_bytes) / k_fused


def ncu_traffic_bytes(workload: str, envs: int, k_fused: int, rng: str = "philox"):
    """DRAM bytes per launch of the step kernel from the committed ncu --set full capture of this
    workload (profiles/r02_ncu_traffic.json); None when no capture matches the configuration."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))[workload + "/" + rng]
        return float(t["dram_bytes_per_launch"]) if (t["envs_per_launch"] == envs and k_fused == 1) else None
    except Exception:
        return None


def pin_rank_to_cores(local_rank: int, local_world: int) -> None:
    """One process per GPU: give each rank its own share of the host cores, so that the pinned staging buffers it allocates
    afterwards and the threads that drive its copies stay put (the box's ranks otherwise pile onto the same cores)."""
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // max(1, local_world))
        mine = cores[local_rank * per:(local_rank + 1) * per] or cores
        os.sched_setaffinity(0, mine)
    except (AttributeError, OSError):
        pass


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm), "reasons": sorted(reasons)}


def cycle_actions(torch, n, dev, real, seed):
    """Uniform actions in the env's Box [0,1] x [0,1] x [-1,1] (salp_robot_env.py:51-55)."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    a = torch.rand(n, 3, device=dev, dtype=real, generator=g)
    a[:, 2] = a[:, 2] * 2 - 1
    return a


def cpu_cycle_oracle_throughput(budget_s: float, threads: int, n_envs: int):
    """Time oracle/salp_cycle_oracle.c (OpenMP over envs) on a bounded sample of the robot_cycle workload."""
    import numpy as np

    from oracle import salp_cycle_oracle
    from underwater_swimmer_rl_b200 import rng_tables
    os.environ["OMP_NUM_THREADS"] = str(threads)
    salp_cycle_oracle.build()
    words = rng_tables.fill_words(n_envs, 64, 10_000, 0, rng_tables.JITTER_STREAM)
    orc = salp_cycle_oracle.CycleOracle(n_envs, words)
    orc.reset()
    g = np.random.default_rng(1234)
    mk = lambda: np.stack([g.uniform(0, 1, n_envs), g.uniform(0, 1, n_envs), g.uniform(-1, 1, n_envs)], axis=1)  # noqa: E731
    orc.step(mk())
    steps = subs = 0
    t0 = time.perf_counter()
    while True:
        subs += orc.step(mk())["substeps"]
        steps += n_envs
        el = time.perf_counter() - t0
        if el >= budget_s:
            break
    return steps / el, subs / el, el, f"{n_envs} envs x {steps // n_envs} cycles, auto-reset, actions uniform in the Box"


def run_cycle_reference_arm(args, wl):
    """--impl reference for robot_cycle: the C restatement of model (B) on all host threads."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = os.cpu_count() or 1
    n = 256 * threads
    v, subs, el, sample = cpu_cycle_oracle_throughput(max(5.0, min(60.0, 0.02 * args.steps)), threads, n)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "envs_per_step": n, "substeps_per_s": subs,
                   "note": "C restatement of the Python reference's model (B) (oracle/salp_cycle_oracle.c), OpenMP over envs"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_cycle_workload(args, wl):
    """robot_cycle_256k: SalpCycleVectorEnv.step (salp_cycle_step_kernel), one breathing cycle per env and step."""
    import torch
    import torch.distributed as dist

    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N = wl["envs"]
    real = torch.float64 if wl["dtype"] == "float64" else torch.float32
    env = SalpCycleVectorEnv(N, device=dev, dtype=wl["dtype"], seed=10_000, env_id0=rank * N, table_rows=min(N, 1 << 16), table_len=64)
    env.reset()
    pool = [cycle_actions(torch, N, dev, real, 1234 + 100 * rank + i) for i in range(8)]
    steps = max(8, min(args.steps, 200))   # one step is milliseconds of device time
    warmup = max(3, min(args.warmup, 8))
    for i in range(warmup):
        env.step(pool[i % 8])
    sampler = ClockSampler(local)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    env.episode_stats(reset=True)
    if rank == 0:
        sampler.start()
    launches0 = env.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(steps):
        env.step(pool[(warmup + i) % 8])
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    launches = env.launch_count - launches0
    st = env.episode_stats()   # NCCL all-reduce (SUM) of the 8-double stats vector when world > 1
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * N * steps / (ms_max * 1e-3)

    rb = 8 if real == torch.float64 else 4
    a_h = cycle_actions(torch, N, dev, real, 99).cpu().pin_memory()
    o_h = torch.empty(N, 6, dtype=torch.float32).pin_memory()
    r_h = torch.empty(N, dtype=real).pin_memory()
    te_h = torch.empty(N, dtype=torch.uint8).pin_memory()
    tr_h = torch.empty(N, dtype=torch.uint8).pin_memory()
    for _ in range(2):
        env.step_host(a_h, o_h, r_h, te_h, tr_h)
    if world > 1:
        dist.barrier()
    e2e_steps = max(4, min(args.e2e_steps, 20))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        env.step_host(a_h, o_h, r_h, te_h, tr_h)
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * N * e2e_steps / float(t.item())

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        # state 6 records + ctrl read and written, actions, obs, reward, two flags, n_sub
        bpe = 2 * (6 * 4 * rb + 16) + 3 * rb + 24 + rb + 2 + 4
        kernel_ms = ms_max / steps
        achieved = bpe * N / (kernel_ms * 1e-3) / 1e9
        sub_per_s = st["substeps"] / (ms_max * 1e-3)
        issue = None
        try:  # warp instructions per substep from the committed ncu capture of this kernel
            prof = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json"))).get(args.workload)
            if prof and clocks and clocks.get("sm_mhz"):
                wi = prof["warp_inst_per_substep"] * sub_per_s / world
                pk = 148 * 4 * clocks["sm_mhz"] * 1e6
                issue = {"warp_inst_per_s": wi, "peak": pk, "frac": wi / pk, "warp_inst_per_substep": prof["warp_inst_per_substep"],
                         "source": "profiles/r01_ncu_traffic.json (smsp__inst_executed.sum / substeps of the captured launch); peak = 148 SMs x 4 schedulers x SM clock"}
        except (OSError, KeyError, ValueError):
            issue = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if real == torch.float64 else "f32", "data": "synthetic",
            "config": {"workload": args.workload, "model": "B (robot.py Robot/Nozzle, salp_robot_env.py SalpRobotEnv)",
                       "envs_per_gpu": N, "global_envs": world * N, "actions": "uniform in Box([0,0,-1],[1,1,1]), 8 resident tensors",
                       "substeps_per_env_step": st["substeps"] / max(st["env_steps"], 1.0), "integrator_substeps_per_s": sub_per_s,
                       "l2": "working set 64 MB (L2-resident); the kernel is ALU-bound, cache state is immaterial",
                       "parallelism": f"env-slab x{world} (no data-path collective; NCCL all-reduce of episode stats only)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "kernel": "salp_cycle_step_kernel", "kernel_ms": kernel_ms, "algorithmic_bytes_per_env_step": bpe,
                         "peak_source": peak_src, "issue": issue,
                         "note": "~1e5 flops per env step against ~300 B of state: the bound is the FP32/SFU issue rate, see issue"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": N * 3 * rb, "d2h_bytes_per_step": N * (24 + rb + 2),
                    "steps": e2e_steps, "api": "SalpCycleVectorEnv.step_host (pinned host buffers) -> salp_cycle_step"},
            "gpu_launches": launches, "clocks": clocks,
            "episode_stats": {k: st[k] for k in ("episodes", "targets_reached", "truncations", "nonfinite", "env_steps", "substeps")},
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, subs, secs, sample = cpu_cycle_oracle_throughput(args.cpu_seconds, threads, 256 * threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample, "seconds": secs,
                                    "substeps_per_s": subs, "note": "C restatement of the Python reference (oracle/salp_cycle_oracle.c), "
                                    "OpenMP; the unmodified Python reference runs ~2.4e3 integrator substeps/s/core in the build container"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def oracle_steady_state(params: dict, n_envs: int, threads: int, prime_steps: int):
    """The CPU arm's batch: oracle envs on the reference's own word streams (MT19937 tables filled inside oracle/ -- the
    product library is never loaded here), primed to the same steady state as the GPU arm (episodes ended, breathing phases
    de-synchronised: resets, respawns and thrust steps inside the timed region).  Returns (Oracle, action generator)."""
    import numpy as np

    from oracle import salp_oracle
    os.environ["OMP_NUM_THREADS"] = str(threads)
    salp_oracle.build()
    fw = salp_oracle.fill_words(n_envs, 128, 10_000, 0, 0)
    jw = salp_oracle.fill_words(n_envs, 512, 10_000, 0, 1)
    orc = salp_oracle.Oracle(n_envs, fw, jw, **params)
    orc.reset()
    g = np.random.default_rng(1234)
    chunk = 256
    for _ in range((prime_steps + chunk - 1) // chunk):
        orc.rollout(g.uniform(-1, 1, (chunk, n_envs, 1)).astype(np.float32), autoreset=True, record=False)
    return orc, g


def cpu_oracle_throughput(params: dict, budget_s: float, threads: int, n_envs: int, T: int, prime_steps: int = 2048):
    """Time the CPU restatement (oracle/salp_oracle.c, OpenMP over envs) on a bounded sample of the same workload in the
    same steady state.  Returns (env-steps/s, seconds, description)."""
    import numpy as np
    orc, g = oracle_steady_state(params, n_envs, threads, prime_steps)
    acts = g.uniform(-1, 1, (T, n_envs, 1)).astype(np.float32)
    done_steps, reps, t0 = 0, 0, time.perf_counter()
    while True:
        orc.rollout(acts, autoreset=True, record=False)
        done_steps += n_envs * T
        reps += 1
        el = time.perf_counter() - t0
        if el >= budget_s or reps >= 64:
            break
    return done_steps / el, el, (f"{n_envs} envs x {T * reps} steps after {prime_steps} priming steps, auto-reset, "
                                 "uniform(-1,1) actions, no output recording")


def cpu_oracle_single_env(params: dict, T: int = 1000, reps: int = 20):
    """BASELINE.json configs[0] with the C restatement: 1 env x 1000 steps on 1 core (best of `reps`)."""
    import numpy as np

    from oracle import salp_oracle
    os.environ["OMP_NUM_THREADS"] = "1"
    fw = salp_oracle.fill_words(1, 128, 10_000, 0, 0)
    jw = salp_oracle.fill_words(1, 512, 10_000, 0, 1)
    acts = np.random.default_rng(1234).uniform(-1, 1, (T, 1, 1)).astype(np.float32)
    best = 0.0
    for _ in range(reps):
        orc = salp_oracle.Oracle(1, fw, jw, **params)
        orc.reset()
        t0 = time.perf_counter()
        orc.rollout(acts, autoreset=True, record=False)
        best = max(best, T / (time.perf_counter() - t0))
    return best


def python_reference_numbers(workload_yaml):
    """The unmodified Python reference timed in the build container (tools/time_python_reference.py ->
    profiles/r02_python_reference_cpu.json; /root/reference cannot travel to the GPU box)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_python_reference_cpu.json")))
        key = os.path.splitext(os.path.basename(workload_yaml or "configs/single_food.yaml"))[0]
        c = d["configs"][key]
        return {"steps_per_s_1core_1env_1000steps": c["steps_per_s_1core_1env_1000steps"], "processes": c["processes"],
                "steps_per_s_all_cores": c["steps_per_s_all_cores_wall"], "machine": "build container (%d cores)" % d["machine"]["cpu_count"],
                "source": "profiles/r02_python_reference_cpu.json (tools/time_python_reference.py)"}
    except (OSError, KeyError, ValueError):
        return None


def run_reference_arm(args, wl):
    """--impl reference: the reference's CPU implementation of the path on the host cores.
    The reference is pure Python and cannot travel to the GPU box, so this is its C restatement
    (oracle/, bit-exact against the Python reference -- DESIGN.md) with all host threads.  Nothing of the product
    package is imported: the word tables come from oracle/ too."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    threads = os.cpu_count() or 1
    n = min(wl["envs"], 1 << 18)
    orc, g = oracle_steady_state(wl["params"], n, threads, args.prime_steps)
    acts = g.uniform(-1, 1, (args.warmup + args.steps, n, 1)).astype(np.float32)
    if args.warmup:
        orc.rollout(acts[: args.warmup], autoreset=True, record=False)
    t0 = time.perf_counter()
    orc.rollout(acts[args.warmup:], autoreset=True, record=False)
    el = time.perf_counter() - t0
    value = n * args.steps / el
    sample = (f"{n} envs x {args.steps} steps per run (bounded sample of {wl['envs']} envs/GPU) after {args.prime_steps} priming steps "
              "(same steady state as the GPU arm), auto-reset")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "reference_yaml": wl["yaml"], "envs_per_step": n, "prime_steps": args.prime_steps,
                   "note": "C restatement of the Python reference (bit-exact, oracle/), OpenMP over envs; the product library is not loaded"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "python_reference": python_reference_numbers(wl["yaml"]),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


class SlabTimer:
    """One env slab of this rank, its resident action pool and a step runner that replays CUDA graphs of up to 32 steps
    (a step of a small slab is shorter than the Python/ctypes launch path; for large slabs the graph keeps the slowest
    rank's launch cadence out of the max-over-ranks time) and runs EXACTLY the requested number of steps."""

    def __init__(self, torch, params, N, dev, dtype, rank_id0, rng, K=1, use_graph=True, seed=1234):
        from underwater_swimmer_rl_b200 import SalpVectorEnv
        self.torch, self.N, self.K = torch, N, K
        real = torch.float64 if dtype == "float64" else torch.float32
        self.env = SalpVectorEnv(N, device=dev, dtype=dtype, seed=10_000, env_id0=rank_id0, food_table_len=128, jitter_table_len=512,
                                 max_substeps=K, autoreset=True, rng=rng, **params)
        self.env.reset()
        gen = torch.Generator(device=dev)
        gen.manual_seed(seed)
        self.pool = [torch.rand(K, N, self.env.act_dim, device=dev, dtype=real, generator=gen) * 2 - 1 for _ in range(32)]
        self.use_graph = use_graph
        self._graphs = {}
        self.i = 0

    def _graph(self, g):
        if g not in self._graphs:
            self._graphs[g] = self.env.capture_graph([p if self.K > 1 else p[0] for p in self.pool[:g]], self.K)
            self._graphs[g]()  # the first launch of a captured graph uploads it to the device (~150 us): never inside a timed region
        return self._graphs[g]

    def prepare(self, n):
        """capture the graphs run(n) will need (outside any timed region)"""
        if self.use_graph:
            if n >= 32:
                self._graph(32)
            if n % 32:
                self._graph(n % 32)

    def run(self, n):
        if not self.use_graph:
            for _ in range(n):
                p = self.pool[self.i % 32]
                self.env.step(p[0]) if self.K == 1 else self.env.step_many(p, self.K)
                self.i += 1
            return
        for _ in range(n // 32):
            self._graph(32)()
        if n % 32:
            self._graph(n % 32)()

    def time(self, steps, warmup, prime, dist=None, world=1):
        """(ms for `steps` steps on this rank, launches) -- CUDA events on the launching stream, barrier + synchronize on both sides"""
        torch = self.torch
        self.prepare(steps), self.prepare(warmup), self.prepare(prime)
        self.run(prime)
        self.run(warmup)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = self.env.launch_count
        ev0.record()
        self.run(steps)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        return ms, (steps if self.use_graph else self.env.launch_count - l0)

    def close(self):
        self._graphs.clear()
        self.env.close()


def rank_max_min(torch, dist, dev, x, world):
    t = torch.tensor([x, -x], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0].item()), -float(t[1].item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="single_food_1m", choices=sorted(WORKLOADS))
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--substeps", type=int, default=1, help="K env steps fused per launch (open-loop actions)")
    ap.add_argument("--rng", default="philox", choices=["philox", "table"],
                    help="philox: both random streams computed in the kernel (unbounded, the mode long rollouts use); table: host-seeded "
                         "MT19937 word tables shared with the reference (the parity mode; finite, cursors wrap)")
    ap.add_argument("--dtype", default="", choices=["", "float32", "float64"], help="override the workload's build (float64 = validation build)")
    ap.add_argument("--torch-actor", action="store_true", help="sac_rollout: torch bf16 MLP instead of the fused tcgen05 actor kernel")
    ap.add_argument("--append-kernel", action="store_true", help="sac_rollout: transition-major DeviceReplayBuffer filled by salp_replay_append "
                    "instead of the zero-copy StepRingReplay (the env's own output rows are the replay storage)")
    ap.add_argument("--graph", default="on", choices=["on", "off"], help="replay the steps from CUDA graphs of up to 32 steps")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary measurements (other RNG mode, strong-scaling block)")
    ap.add_argument("--prime-steps", type=int, default=2048,
                    help="untimed env steps before warm-up so episodes have ended and envs are de-synchronised "
                         "(steady state: resets, respawns and divergent breathing phases inside the timed region)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    wl = dict(WORKLOADS[args.workload])
    if args.envs_per_gpu:
        wl["envs"] = args.envs_per_gpu
    if args.dtype:
        wl["dtype"] = args.dtype
    rollout = bool(wl.get("rollout"))
    spr = 1
    if rollout:
        spr = 2 if args.append_kernel else RING_STEPS  # env steps per rollout graph replay
        args.steps = (args.steps + spr - 1) // spr * spr
        args.warmup = (args.warmup + spr - 1) // spr * spr
    if wl.get("cycle"):
        return run_cycle_reference_arm(args, wl) if args.impl == "reference" else run_cycle_workload(args, wl)
    if args.impl == "reference":
        return run_reference_arm(args, wl)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    pin_rank_to_cores(local, int(os.environ.get("LOCAL_WORLD_SIZE", str(world))))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    N, K = wl["envs"], (1 if rollout else args.substeps)
    real = torch.float64 if wl["dtype"] == "float64" else torch.float32
    rb = 8 if real == torch.float64 else 4
    if rollout:
        slab = None
        from underwater_swimmer_rl_b200 import SalpVectorEnv
        env = SalpVectorEnv(N, device=dev, dtype=wl["dtype"], seed=10_000, env_id0=rank * N, food_table_len=128, jitter_table_len=512,
                            max_substeps=2 if args.append_kernel else RING_STEPS, autoreset=True, rng=args.rng, **wl["params"])
        env.reset()
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234 + rank)
        pool = [torch.rand(1, N, env.act_dim, device=dev, dtype=real, generator=gen) * 2 - 1 for _ in range(2)]
        from underwater_swimmer_rl_b200.rollout import (DeviceReplayBuffer, FusedActor, RolloutCollector, StepRingReplay,
                                                         TanhGaussianActor)
        torch.manual_seed(1234 + rank)
        actor = TanhGaussianActor(env.obs_dim, env.act_dim, (256, 256)).to(dev)
        actor = actor.to(torch.bfloat16) if args.torch_actor else FusedActor(actor)
        ring = DeviceReplayBuffer(4 * N, env.obs_dim, env.act_dim, dev) if args.append_kernel else StepRingReplay(env)
        col = RolloutCollector(env, actor, ring)
        replay = col.capture()  # one graph launch = spr env steps of (actor -> env kernel [-> replay append])
        assert col.steps_per_replay == spr
        launches_per_step = (0 if args.torch_actor else 1) + 1 + (1 if args.append_kernel else 0)

        def run(n):
            for _ in range(n // spr):
                replay()
        run((args.prime_steps + spr - 1) // spr * spr)
        run(args.warmup)
    else:
        slab = SlabTimer(torch, wl["params"], N, dev, wl["dtype"], rank * N, args.rng, K, args.graph == "on", 1234 + rank)
        env = slab.env

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if rollout:
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        run(args.steps)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        launches = launches_per_step * args.steps
    else:
        ms, launches = slab.time(args.steps, args.warmup, args.prime_steps, dist, world)
    clocks = sampler.stop() if rank == 0 else None
    ms_max, ms_min = rank_max_min(torch, dist, dev, ms, world)
    value = world * N * K * args.steps / (ms_max * 1e-3)

    env_only_ms = None
    if rollout:  # the env kernel alone, for the roofline of THE kernel inside the rollout step
        env_replay = env.capture_graph(pool[0][0])  # graph replay: the Python launch path is slower than this kernel
        for i in range(10):
            env_replay()
        torch.cuda.synchronize()
        ev0.record()
        for i in range(200):
            env_replay()
        ev1.record()
        torch.cuda.synchronize()
        env_only_ms = ev0.elapsed_time(ev1) / 200

    # ---- e2e: the reference-facing call with HOST buffers (H2D actions, step, D2H results) ----
    a_h = (torch.rand(N, env.act_dim, dtype=real) * 2 - 1).pin_memory()
    o_h = torch.empty(N, env.obs_dim, dtype=torch.float32).pin_memory()
    r_h = torch.empty(N, dtype=real).pin_memory()
    te_h = torch.empty(N, dtype=torch.bool).pin_memory()
    tr_h = torch.empty(N, dtype=torch.bool).pin_memory()
    for _ in range(3):
        env.step_host(a_h, o_h, r_h, te_h, tr_h)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        env.step_host(a_h, o_h, r_h, te_h, tr_h)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_max, e2e_min = rank_max_min(torch, dist, dev, e2e_s, world)
    e2e_value = world * N * args.e2e_steps / e2e_max
    h2d = N * env.act_dim * rb
    d2h = N * (env.obs_dim * 4 + rb + 2)
    # the ceiling the box puts on that call: every rank copies its observation rows device -> pinned host memory at the same
    # time, no kernel (what the PCIe links / switches and the host memory sustain with `world` GPUs copying together)
    obs_dev = env.obs[0]
    for _ in range(2):
        o_h.copy_(obs_dev, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(10):
        o_h.copy_(obs_dev, non_blocking=True)
    torch.cuda.synchronize()
    copy_s = time.perf_counter() - t0
    copy_max, _ = rank_max_min(torch, dist, dev, copy_s, world)
    d2h_copy_only_gbs = 10 * obs_dev.numel() * 4 / copy_max / 1e9

    stats = env.episode_stats()  # NCCL all-reduce (SUM) of the 8-double stats vector when world > 1
    obs_dim, act_dim = env.obs_dim, env.act_dim
    peak, peak_src = measured_peak_gbs()
    nf = wl["params"]["num_food_items"]

    # ---- secondary measurements (same protocol, every rank, max over ranks): the other RNG mode at the same slab size, and
    #      the literal north_star split -- 1 048 576 envs over the GPUs of the job, for single_food and defaults
    extras = {}
    if not rollout and not args.no_extras and K == 1:
        slab.close()
        del slab, env
        torch.cuda.empty_cache()
        other = "table" if args.rng == "philox" else "philox"
        s2 = SlabTimer(torch, wl["params"], N, dev, wl["dtype"], rank * N, other, 1, args.graph == "on", 1234 + rank)
        ms2, _ = s2.time(max(args.steps, 64), args.warmup, args.prime_steps, dist, world)
        ms2_max, _ = rank_max_min(torch, dist, dev, ms2 / max(args.steps, 64), world)
        s2.close()
        del s2
        torch.cuda.empty_cache()
        extras["rng_modes"] = {args.rng: {"ms_per_step": ms_max / args.steps, "value": value},
                               other: {"ms_per_step": ms2_max, "value": world * N / (ms2_max * 1e-3)},
                               "note": "philox: Philox4x32-10 in the kernel, keyed by (seed, global env id, word index): unbounded streams, no "
                                       "memory; table: host-seeded MT19937 words shared with the reference (parity mode, 128 + 512 words per env)"}
        n_s = (1 << 20) // world
        strong = {"global_envs": n_s * world, "envs_per_gpu": n_s, "rng": args.rng, "steps": max(args.steps, 64)}
        for name, prm in (("single_food", SINGLE_FOOD), ("defaults", DEFAULTS)):
            st_ = SlabTimer(torch, prm, n_s, dev, "float32", rank * n_s, args.rng, 1, args.graph == "on", 1234 + rank)
            ms3, _ = st_.time(max(args.steps, 64), args.warmup, args.prime_steps, dist, world)
            ms3_max, _ = rank_max_min(torch, dist, dev, ms3 / max(args.steps, 64), world)
            st_.close()
            del st_
            torch.cuda.empty_cache()
            b3 = algorithmic_bytes_per_env_step(prm["num_food_items"], 4, 1, 1, 24)
            ach = b3 * n_s / (ms3_max * 1e-3) / 1e9
            strong[name] = {"ms_per_step_per_gpu": ms3_max, "value": n_s * world / (ms3_max * 1e-3), "achieved_gbs_per_gpu": ach,
                            "frac": ach / peak, "algorithmic_bytes_per_env_step": b3}
        extras["strong"] = strong
        if wl["dtype"] == "float32":
            # the fp64 validation build (bit-exact discrete trajectories against the reference) on the same workload,
            # same protocol: north_star couples ">= 60 % of the HBM roofline" with bit-exact discrete state
            s4 = SlabTimer(torch, wl["params"], N, dev, "float64", rank * N, args.rng, 1, args.graph == "on", 1234 + rank)
            n4 = max(min(args.steps, 400), 64)
            ms4, _ = s4.time(n4, args.warmup, args.prime_steps, dist, world)
            ms4_max, _ = rank_max_min(torch, dist, dev, ms4 / n4, world)
            s4.close()
            del s4
            torch.cuda.empty_cache()
            b4 = algorithmic_bytes_per_env_step(nf, 8, 1, act_dim, obs_dim)
            ach4 = b4 * N / (ms4_max * 1e-3) / 1e9
            extras["fp64_validation_build"] = {"ms_per_step": ms4_max, "value": world * N / (ms4_max * 1e-3), "steps": n4,
                                               "algorithmic_bytes_per_env_step": b4, "achieved_gbs_per_gpu": ach4, "frac": ach4 / peak,
                                               "note": "dtype f64, reference operation order, discrete trajectories bit-exact vs the oracle "
                                                       "(tests/test_gpu_parity.py); same workload, slab size and timing protocol as `value`"}

    if rank == 0:
        bpe = algorithmic_bytes_per_env_step(nf, rb, K, act_dim, obs_dim)
        bmin = min_traffic_bytes_per_env_step(nf, rb, K, act_dim, obs_dim)
        kernel_ms = ms_max / args.steps  # the timed region is nothing but back-to-back launches of this kernel
        if rollout:
            kernel_ms = env_only_ms  # env kernel timed alone; the step also holds the actor GEMMs and the append
        achieved = bpe * N * K / (kernel_ms * 1e-3) / 1e9
        fast = (real == torch.float32 and wl["params"].get("forced_breathing", True) and nf in (1, 12))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if real == torch.float64 else "f32", "data": "synthetic",
            "config": {"workload": args.workload, "reference_yaml": wl["yaml"], "envs_per_gpu": N, "global_envs": world * N,
                       "substeps_per_launch": K, "rng": args.rng,
                       "cuda_graph": ("graphs of up to 32 steps" if args.graph == "on" else None) if not rollout else ("%d steps per replay" % spr),
                       "rollout": ("actor MLP [24-256-256-2] bf16 (" + ("torch" if args.torch_actor else "salp_actor_kernel, tcgen05") + ") -> salp_step kernel -> " + ("salp_replay_append (transition-major ring)" if args.append_kernel else "zero-copy step ring (the env's [5, N] output rows are the replay storage: a 4-step window per env, not a uniform 1M-transition history; --append-kernel for the transition-major ring)") + ", CUDA graph of %d steps" % spr + "; env kernel share of the step = %.2f" % (env_only_ms / (ms_max / args.steps))) if rollout else None,
                       "prime_steps": args.prime_steps, "obs_dim": obs_dim, "act_dim": act_dim, "autoreset": True,
                       "l2": f"per-step footprint {bpe * N / 1e6:.0f} MB vs 126 MB L2"
                             + (" (inputs larger than L2, no flush needed)" if bpe * N > 126e6 else " (L2-resident working set)"),
                       "parallelism": f"env-slab x{world} (no data-path collective; NCCL all-reduce of episode stats only)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic_bytes(args.workload, N, K, args.rng),
                         "kernel": "salp_step_fast_kernel<%d, %d>" % (nf, 1 if args.rng == "philox" else 0) if fast else "salp_step_kernel",
                         "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_env_step": bpe, "algorithmic_bytes_per_launch": bpe * N * K,
                         "min_traffic_bytes_per_env_step": bmin, "achieved_min_traffic": bmin * N * K / (kernel_ms * 1e-3) / 1e9,
                         "frac_min_traffic": bmin * N * K / (kernel_ms * 1e-3) / 1e9 / peak,
                         "traffic_source": "profiles/r02_ncu_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum per launch)",
                         "peak_source": peak_src},
            "per_rank_ms_per_step": {"min": ms_min / args.steps, "max": ms_max / args.steps},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": args.e2e_steps, "api": "SalpVectorEnv.step_host -> salp_step_host (pinned host buffers)",
                    "d2h_gbs_per_gpu": d2h * args.e2e_steps / e2e_max / 1e9, "per_rank_s": {"min": e2e_min, "max": e2e_max},
                    "d2h_copy_only_gbs_per_gpu": d2h_copy_only_gbs,
                    "d2h_copy_only_note": "all %d ranks copying their observation rows to pinned host memory at once, no kernel: the "
                                          "box's ceiling for this call (slowest rank)" % world},
            "gpu_launches": launches, "clocks": clocks,
            "episode_stats": {k: stats[k] for k in ("episodes", "mean_return", "mean_length", "collisions", "truncations", "env_steps")},
        }
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, secs, sample = cpu_oracle_throughput(wl["params"], args.cpu_seconds, threads, 4096 * max(1, min(threads, 16)), 250, args.prime_steps)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                                    "seconds": secs, "note": "C restatement of the Python reference (oracle/salp_oracle.c), OpenMP over envs"}
            line["cpu_baseline_1core"] = {"value": cpu_oracle_single_env(SINGLE_FOOD), "unit": UNIT, "cores": 1, "kind": "port",
                                          "sample": "BASELINE configs[0]: single_food.yaml, 1 env x 1000 steps, best of 20"}
            line["python_reference"] = python_reference_numbers(wl["yaml"])
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
