"""Time the UNMODIFIED Python reference (SalpSnakeEnv over the legacy SalpRobotEnv, /root/reference) on this machine's
host cores -- BASELINE.json configs[0] and the BASELINE.md section 3 CPU plan.  Build container only (the reference tree
cannot travel to the GPU box); writes profiles/r02_python_reference_cpu.json, which bench.py quotes next to its own
cpu_baseline (the C restatement timed on the GPU box).

    python tools/time_python_reference.py

  (i)  1 core: 1 env x 1000 steps (configs[0]) and 1 env x 20000 steps, seeded uniform(-1, 1) actions, reset on done
  (ii) all cores: P = os.cpu_count() processes, one env each, aggregate steps/s
for the four reference YAMLs' environment.params.
"""
import json
import multiprocessing as mp
import os
import platform
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402
from oracle.validate_against_reference import CONFIGS  # noqa: E402


def run_loop(params, seed, T):
    """The caller's loop: for t in range(T): env.step(a[t]); reset on done/truncated."""
    env = rh.make_reference_env(**params)
    rh.seed_reference(seed)
    acts = np.random.default_rng(1234 + seed).uniform(-1, 1, (T, 1)).astype(np.float32)
    env.reset()
    t0 = time.perf_counter()
    for t in range(T):
        _, _, d, tr, _ = env.step(acts[t])
        if d or tr:
            env.reset()
    return T / (time.perf_counter() - t0)


def _worker(args):
    params, seed, T = args
    return run_loop(params, seed, T)


def main():
    assert rh.reference_available(), "needs /root/reference"
    P = os.cpu_count() or 1
    out = {"what": "unmodified Python reference (scripts/utilities/salp_robot.py + src/salp/environments/salp_snake_env.py), "
                   "render_mode=None, uniform(-1,1) float32 actions, reset on done; time.perf_counter around the step loop",
           "machine": {"cpu_count": P, "python": platform.python_version(), "numpy": np.__version__, "platform": platform.platform(),
                       "note": "build container, not the GPU box"},
           "configs": {}}
    for name, params in CONFIGS.items():
        one_1k = max(run_loop(params, 10_000, 1000) for _ in range(3))
        one_20k = run_loop(params, 10_001, 20_000)
        with mp.Pool(P) as pool:
            t0 = time.perf_counter()
            rates = pool.map(_worker, [(params, 20_000 + i, 10_000) for i in range(P)])
            wall = time.perf_counter() - t0
        out["configs"][name] = {"steps_per_s_1core_1env_1000steps": one_1k, "steps_per_s_1core_1env_20000steps": one_20k,
                                "processes": P, "steps_per_s_all_cores_sum_of_rates": float(sum(rates)),
                                "steps_per_s_all_cores_wall": P * 10_000 / wall}
        print(name, out["configs"][name], flush=True)
    path = os.path.join(ROOT, "profiles", "r02_python_reference_cpu.json")
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
